"""One process, two GPUs (engine.cu:778): per-engine head scratch on the engine's own device and per-device
`cudaFuncSetAttribute` opt-ins.  Skipped on single-GPU boxes; run under `gpurun --gpus 2`."""

import pytest
import torch

from deepfakedetection_b200 import _lib, registry
from deepfakedetection_b200.gradcam import ClassifierOutputTarget, GradCAM

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,oracle_mod", [("efficientnet_b0", "oracle.effnet_b0"), ("efficientformerv2_s1", "oracle.efficientformerv2"),
                                             ("faster_vit_0_224", "oracle.fastervit")])
def test_forward_and_hook_gradcam_on_cuda0_then_cuda1(name, oracle_mod):
    import importlib

    from oracle.effnet_b0 import synth_images

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    oracle = importlib.import_module(oracle_mod).make_seeded_oracle(seed=0)
    x = synth_images(4, seed=2)
    with torch.no_grad():
        ref = oracle(x)
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        model = registry.get_model_spec(name).builder(name, 2)
        model.to(torch.device(dev)).eval()
        model.load_state_dict(oracle.state_dict(), strict=False)
        with torch.inference_mode():
            out = model(x.to(dev))
        assert out.device == torch.device(dev)
        outs.append(out.cpu())
        assert (out.cpu().softmax(1) - ref.softmax(1)).abs().max().item() <= 2e-2
        if model.supports_cam_hooks:   # the hook protocol allocates the per-engine head scratch on THIS device
            with GradCAM(model=model, target_layers=[model._cam_point()], use_hooks=True) as cam:
                gray = cam(input_tensor=x[:1].to(dev), targets=[ClassifierOutputTarget(int(ref[0].argmax()))])[0]
            assert gray.shape == (224, 224) and 0.0 <= float(gray.min()) and float(gray.max()) <= 1.0 + 1e-6
    assert torch.equal(outs[0], outs[1]), "the same weights and inputs must give the same bits on both devices"
    # the same module moved across devices rebuilds its engine there
    model.to("cuda:0")
    with torch.inference_mode():
        again = model(x.to("cuda:0")).cpu()
    assert torch.equal(again, outs[0])
