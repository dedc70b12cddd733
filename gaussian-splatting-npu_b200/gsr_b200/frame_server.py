"""Headless frame server: the TCP transport of the reference's network viewer hook, without the viewer.

Mirrors /root/reference/gaussian_renderer/network_gui.py:26-85 -- same wire protocol, same call names -- on top of
`frames.camera_from_message` / `frames.frame_bytes`:

  request  (viewer -> trainer):  u32 little-endian length, then that many bytes of UTF-8 JSON
                                 (resolution_x/y, train, fov_x/y, z_near/far, shs_python, rot_scale_python, keep_alive,
                                  scaling_modifier, view_matrix[16], view_projection_matrix[16])     network_gui.py:43-85
  response (trainer -> viewer):  [H*W*3 bytes of RGB8, only when a frame was rendered], u32 length, `verify` in ASCII
                                 (the dataset path the viewer checks)                                network_gui.py:49-53

`FrameServer` keeps the reference's surface (`init`, `try_connect`, `read`, `send`, `receive`, `.conn`) so the loop of
train.py:74-87 reads the same; `serve_pending(render, verify)` is that loop body.  Differences from the reference: the
listener is created by `init` (not at import), reads loop until the announced length has arrived (`conn.recv(n)` may
return fewer bytes), and the frame comes from one device launch (gsr_frame_to_rgb8).  Plain sockets: no graphics stack
is involved; the SIBR viewer itself (OpenGL) is unavailable on this box -- any client speaking the protocol works.
"""
import json
import socket

from . import frames


class FrameServer:
    def __init__(self):
        self.host, self.port = "127.0.0.1", 6009
        self.listener = None
        self.conn = None
        self.addr = None

    # ---- network_gui.py:26-41 ----
    def init(self, wish_host="127.0.0.1", wish_port=6009):
        self.host, self.port = wish_host, wish_port
        self.listener = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self.listener.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self.listener.bind((self.host, self.port))
        self.listener.listen()
        self.listener.settimeout(0)
        self.port = self.listener.getsockname()[1]          # wish_port 0: the kernel picked one
        return self.port

    def try_connect(self):
        if self.listener is None:
            return False
        try:
            self.conn, self.addr = self.listener.accept()
            self.conn.settimeout(None)
            return True
        except (BlockingIOError, socket.timeout, OSError):
            return False

    def _recv_exact(self, n):
        buf = bytearray()
        while len(buf) < n:
            part = self.conn.recv(n - len(buf))
            if not part:
                raise ConnectionError("viewer closed the connection")
            buf += part
        return bytes(buf)

    # ---- network_gui.py:43-47 ----
    def read(self):
        length = int.from_bytes(self._recv_exact(4), "little")
        return json.loads(self._recv_exact(length).decode("utf-8"))

    # ---- network_gui.py:49-53 ----
    def send(self, message_bytes, verify):
        if message_bytes is not None:
            self.conn.sendall(message_bytes)
        self.conn.sendall(len(verify).to_bytes(4, "little"))
        self.conn.sendall(bytes(verify, "ascii"))

    # ---- network_gui.py:55-85 ----
    def receive(self):
        return frames.camera_from_message(self.read())

    def close(self):
        for s in (self.conn, self.listener):
            if s is not None:
                try:
                    s.close()
                except OSError:
                    pass
        self.conn = self.listener = None

    def serve_pending(self, render, verify, training_allowed=True):
        """The viewer block of the training loop (train.py:74-87): answer requests until the viewer lets training
        continue.  `render(cam, scaling_modifier, shs_python, rot_scale_python)` returns the [3,H,W] CUDA image.
        Returns the number of frames sent; a broken connection is dropped like the reference does."""
        if self.conn is None:
            self.try_connect()
        sent = 0
        while self.conn is not None:
            try:
                cam, do_training, shs_py, rot_scale_py, keep_alive, scaling_modifier = self.receive()
                payload = None
                if cam is not None:
                    payload = frames.frame_bytes(render(cam, scaling_modifier, shs_py, rot_scale_py))
                    sent += 1
                self.send(payload, verify)
                if do_training and (training_allowed or not keep_alive):
                    break
            except Exception:
                try:
                    self.conn.close()
                except OSError:
                    pass
                self.conn = None
        return sent


_default = FrameServer()
# module-level aliases: `from gsr_b200 import frame_server as network_gui` reads like the reference's module
init, try_connect, read, send, receive = _default.init, _default.try_connect, _default.read, _default.send, _default.receive
