"""Null-object pyplot: every call succeeds and returns another null object; `savefig(path)` leaves a 1x1 PNG so the
reference's run directory has the files it expects."""


class _Null:
    def __call__(self, *args, **kwargs):
        return _Null()

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Null()

    def __iter__(self):
        return iter([_Null()])

    def __getitem__(self, key):
        return _Null()

    def __float__(self):
        return 0.0

    def savefig(self, path, *args, **kwargs):
        from PIL import Image

        Image.new("RGB", (1, 1)).save(str(path), format="PNG")


def subplots(*args, **kwargs):
    return _Null(), _Null()


def close(*args, **kwargs):
    return None


def __getattr__(name):
    if name.startswith("__") and name.endswith("__"):
        raise AttributeError(name)
    return _Null()
