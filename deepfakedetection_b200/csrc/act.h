// Activation ids shared by every kernel's `act` argument (host and device code).
#pragma once

namespace b200df {
enum Act : int { kActNone = 0, kActSiLU = 1, kActGELU = 2, kActReLU = 3 };
}  // namespace b200df
