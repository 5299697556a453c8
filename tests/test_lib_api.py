"""`lib/frame2frame` entry points (frame2frame_b200/lib_api.py): config keys/defaults and error behaviour on CPU,
checkpoint loading and the in-tree forward convention on the GPU."""
import numpy as np
import pytest
import torch


def test_extract_config_defaults_and_errors():
    from frame2frame_b200 import lib_api
    c = lib_api.extract_config({})
    # dncnn/io.py:64-76 defaults
    assert c["channels"] == 3 and c["num_of_layers"] == 17 and c["net_name"] == "dncnn"
    assert c["pretrained_load"] is False and c["pretrained_path"] == "" and c["device"] == "cuda:0"
    c = lib_api.extract_config({"channels": 1, "unrelated_key": 5, "pretrained_load": True})
    assert c["channels"] == 1 and "unrelated_key" not in c and c["pretrained_load"] is True     # restrict=True
    with pytest.raises(ValueError):
        lib_api.extract_config({"net_name": "fastdvdnet"})      # outside the scope of this build
    with pytest.raises(ValueError):
        lib_api.extract_config({"net_name": "nonsense"})        # __init__.py:31,39
    m = lib_api.load_model({"device": "cpu"})                   # the reference's default: channels = 3 (dncnn/io.py:64-67)
    sd = m.state_dict()
    assert tuple(sd["dncnn.0.weight"].shape) == (64, 3, 3, 3) and tuple(sd["dncnn.47.weight"].shape) == (3, 64, 3, 3)
    assert len(sd) == 92 and m.flat_params.numel() == 556032 + 2 * 1152
    with pytest.raises(NotImplementedError):
        lib_api.load_model({"device": "cpu", "channels": 2})   # only gray and RGB are built


@pytest.mark.gpu
def test_load_model_checkpoint_and_forward_convention(cuda, tmp_path):
    from frame2frame_b200 import lib_api, synth
    from oracle import dncnn_oracle as do
    sd = synth.pretrained_state_dict()
    torch.save({"state_dict": {k: torch.from_numpy(v) for k, v in sd.items()}}, str(tmp_path / "ckpt.pth"))
    model = lib_api.load_model({"channels": 1, "device": "cuda:0", "pretrained_load": True,
                                "pretrained_path": "ckpt.pth", "pretrained_root": str(tmp_path)})
    ref = do.load_numpy_state(do.OracleDnCNN(residual=True), sd)
    H, W = 40, 56
    x = torch.from_numpy(np.stack([synth.noisy_frame(H, W, t) / 255.0 for t in range(2)])[None, :, None].astype(np.float32))
    model.eval(); ref.eval()
    with torch.no_grad():
        y = model(x.to(cuda))                                   # 5-D video batch, net.py:25-37
        yr = ref(x)
    assert y.shape == x.shape
    rel = float((y.cpu() - yr).norm() / yr.norm())
    assert rel <= 3e-2                                          # x - dncnn(x), net.py:31
    crit = lib_api.WarpedLoss(dist_crit="l1")                   # warped_loss.py:118-131
    flow = torch.from_numpy(synth.backward_flow(H, W, 1))[None]
    loss = crit(y[0, 1:2], x[0, 0:1].to(cuda), flow, 0)
    assert loss.ndim == 0 and float(loss) > 0


def test_reference_package_name_is_importable():
    """`import frame2frame` -- what an unmodified harness does (setup.py:18-32, lib/frame2frame/__init__.py:1-39)."""
    import frame2frame
    from frame2frame_b200 import lib_api
    assert frame2frame.WarpedLoss is lib_api.WarpedLoss and frame2frame.load_model is lib_api.load_model
    assert frame2frame.extract_model_config({"channels": 1})["channels"] == 1
    assert frame2frame.dncnn.load_model is lib_api.load_model and frame2frame.dncnn.DnCNN.__name__ == "DnCNN"
    with pytest.raises(ImportError):
        frame2frame.lightning                                       # out of scope, says why
    with pytest.raises(AttributeError):
        frame2frame.nonsense
