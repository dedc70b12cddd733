"""Frame output, CPU side: the oracle (oracle/frame_oracle.py) against the reference's torch expressions
(/root/reference/train.py:81, torchvision.utils.save_image as called at /root/reference/render.py:44), the PNG writer
against PIL's decoder and torchvision's writer, and the viewer-request parser against network_gui.py:55-85."""
import io

import numpy as np
import pytest
import torch

from oracle import frame_oracle as fo
from gsr_b200 import frames


def make(H, W, seed):
    rng = np.random.default_rng(seed)
    img = (rng.random((3, H, W)) * 1.4 - 0.2).astype(np.float32)
    k = np.arange(256, dtype=np.float32)
    edge = np.concatenate([k / 255.0, (k + 0.5) / 255.0, np.nextafter(k / 255, 0).astype(np.float32), [0.0, 1.0, -0.0, 2.0, -1.0]])
    n = min(edge.size, H * W)
    img.reshape(3, -1)[:, :n] = edge[:n]                      # values on the quantisation boundaries
    return img


@pytest.mark.parametrize("H,W", [(1, 1), (5, 7), (33, 130)])
def test_oracle_vs_reference_torch_expressions(H, W):
    img = make(H, W, H * W)
    t = torch.from_numpy(img)
    want_gui = (torch.clamp(t, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous().numpy()          # train.py:81
    assert np.array_equal(fo.to_rgb8(img, "truncate"), want_gui)
    want_png = t.clone().mul(255).add_(0.5).clamp_(0, 255).permute(1, 2, 0).to("cpu", torch.uint8).numpy()  # save_image
    assert np.array_equal(fo.to_rgb8(img, "round"), want_png)
    assert np.array_equal(fo.to_rgb8(img, "round", W // 2, W), want_png[:, W // 2:])                        # render.py:40-42


def test_png_writer_vs_pil_and_torchvision(tmp_path):
    PIL = pytest.importorskip("PIL.Image")
    tv = pytest.importorskip("torchvision.utils")
    img = make(37, 53, 1)
    rgb = fo.to_rgb8(img, "round")
    data = frames.encode_png(rgb)
    back = np.array(PIL.open(io.BytesIO(data)).convert("RGB"))
    assert np.array_equal(back, rgb)
    p = str(tmp_path / "ref.png")
    tv.save_image(torch.from_numpy(img), p)                                                                # render.py:44
    assert np.array_equal(np.array(PIL.open(p).convert("RGB")), back)


def test_camera_from_message_vs_network_gui_receive():
    rng = np.random.default_rng(2)
    vm, vpm = rng.standard_normal(16), rng.standard_normal(16)
    msg = dict(resolution_x=320, resolution_y=200, train=1, fov_y=0.6, fov_x=0.9, z_near=0.01, z_far=100.0, shs_python=0,
               rot_scale_python=1, keep_alive=1, scaling_modifier=0.75, view_matrix=[float(v) for v in vm],
               view_projection_matrix=[float(v) for v in vpm])               # json.loads gives Python floats
    cam, do_training, shs_py, rs_py, keep, sm = frames.camera_from_message(msg)
    wv = torch.reshape(torch.tensor(msg["view_matrix"]), (4, 4))          # network_gui.py:72-78 restated with CPU tensors
    wv[:, 1] = -wv[:, 1]
    wv[:, 2] = -wv[:, 2]
    fp = torch.reshape(torch.tensor(msg["view_projection_matrix"]), (4, 4))
    fp[:, 1] = -fp[:, 1]
    assert np.array_equal(cam.world_view_transform, wv.numpy()) and np.array_equal(cam.full_proj_transform, fp.numpy())
    assert np.allclose(cam.camera_center, torch.inverse(wv)[3][:3].numpy(), rtol=1e-4, atol=1e-5)
    assert (do_training, shs_py, rs_py, keep, sm) == (True, False, True, True, 0.75)
    assert (cam.image_width, cam.image_height, cam.FoVy, cam.FoVx) == (320, 200, 0.6, 0.9)
    assert frames.camera_from_message(dict(msg, resolution_x=0)) == (None,) * 6


def test_no_cpu_path():
    with pytest.raises(RuntimeError, match="no CPU path"):
        frames.to_rgb8(torch.zeros(3, 4, 4))
