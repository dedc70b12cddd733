"""Headless frame output: device render -> bytes / PNG, no viewer and no GL.

Host mirror of
  /root/reference/train.py:74-87                       the frame the training loop ships to network_gui: `frame_bytes`
  /root/reference/gaussian_renderer/network_gui.py:55-85  the viewer's camera request -> MiniCam: `camera_from_message`
  /root/reference/render.py:30-46                      render_set: one PNG per view (torchvision.utils.save_image):
                                                       `save_png` / `encode_png`
over include/gsr_frame.h (gsr_frame_to_rgb8 in libgsr_b200.so: one launch per frame).  The PNG container (zlib stream +
chunks) is written on the host with the standard library; the pixels are quantised on the device.  No CPU path for the
conversion.  The socket transport of network_gui is not rebuilt (networking is outside the hot-path scope)."""
import ctypes
import struct
import zlib

import numpy as np
import torch

from . import capi

TRUNCATE, ROUND = 0, 1
_bound = False


def _lib():
    global _bound
    L = capi.lib()
    if not _bound:
        L.gsr_frame_to_rgb8.restype = ctypes.c_int
        L.gsr_frame_to_rgb8.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p] * 3
        _bound = True
    return L


def to_rgb8(image, mode=TRUNCATE, x_begin=0, x_end=None, out=None):
    """[3,H,W] fp32 CUDA tensor -> [H, x_end-x_begin, 3] uint8 CUDA tensor (gsr_frame_to_rgb8)."""
    if not image.is_cuda:
        raise RuntimeError("to_rgb8 needs a CUDA tensor: there is no CPU path")
    assert image.dim() == 3 and image.shape[0] == 3 and image.dtype == torch.float32
    image = image.contiguous()
    H, W = image.shape[1:]
    x_end = W if x_end is None else x_end
    if out is None:
        out = torch.empty(H, max(x_end - x_begin, 0), 3, dtype=torch.uint8, device=image.device)
    else:
        assert out.is_cuda and out.dtype == torch.uint8 and out.is_contiguous() and out.numel() == H * max(x_end - x_begin, 0) * 3
    with torch.cuda.device(image.device):
        capi._check(_lib().gsr_frame_to_rgb8(H, W, x_begin, x_end, mode, capi._ptr(image), capi._ptr(out), capi._stream()))
    return out


def frame_bytes(net_image):
    """train.py:81: the memoryview network_gui.send() writes to the socket (H*W*3 bytes, row-major RGB)."""
    return memoryview(to_rgb8(net_image, TRUNCATE).cpu().numpy())


def encode_png(rgb8, compress_level=6):
    """[H,W,3] uint8 (numpy or tensor) -> PNG file bytes: 8-bit RGB, filter type 0 on every row."""
    a = rgb8.cpu().numpy() if isinstance(rgb8, torch.Tensor) else np.asarray(rgb8)
    assert a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] == 3
    H, W, _ = a.shape
    rows = np.zeros((H, 1 + 3 * W), np.uint8)
    rows[:, 1:] = a.reshape(H, 3 * W)

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data))
    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", W, H, 8, 2, 0, 0, 0))
            + chunk(b"IDAT", zlib.compress(rows.tobytes(), compress_level)) + chunk(b"IEND", b""))


def save_png(rendering, path, right_half=False):
    """render.py:40-45: torchvision.utils.save_image(rendering, path) for one [3,H,W] render (right_half: the
    train_test_exp crop `rendering[..., W // 2:]`)."""
    W = rendering.shape[-1]
    data = encode_png(to_rgb8(rendering, ROUND, W // 2 if right_half else 0, W))
    with open(path, "wb") as f:
        f.write(data)


class MiniCam:
    """scene/cameras.py:92-103 with numpy arrays (row-vector convention, as the rasterizer settings take them)."""

    def __init__(self, width, height, fovy, fovx, znear, zfar, world_view_transform, full_proj_transform):
        self.image_width, self.image_height = width, height
        self.FoVy, self.FoVx, self.znear, self.zfar = fovy, fovx, znear, zfar
        self.world_view_transform = world_view_transform
        self.full_proj_transform = full_proj_transform
        self.camera_center = np.linalg.inv(world_view_transform.astype(np.float64))[3, :3].astype(np.float32)
        self.tanfovx, self.tanfovy = float(np.tan(fovx * 0.5)), float(np.tan(fovy * 0.5))


def camera_from_message(message):
    """network_gui.receive (network_gui.py:55-85): the viewer's JSON request -> (MiniCam, do_training, do_shs_python,
    do_rot_scale_python, keep_alive, scaling_modifier), or six Nones for a zero-sized request."""
    width, height = message["resolution_x"], message["resolution_y"]
    if width == 0 or height == 0:
        return None, None, None, None, None, None
    wv = np.array(message["view_matrix"], np.float32).reshape(4, 4)
    wv[:, 1] = -wv[:, 1]
    wv[:, 2] = -wv[:, 2]
    fp = np.array(message["view_projection_matrix"], np.float32).reshape(4, 4)
    fp[:, 1] = -fp[:, 1]
    cam = MiniCam(width, height, message["fov_y"], message["fov_x"], message["z_near"], message["z_far"], wv, fp)
    return (cam, bool(message["train"]), bool(message["shs_python"]), bool(message["rot_scale_python"]),
            bool(message["keep_alive"]), message["scaling_modifier"])
