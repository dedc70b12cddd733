"""The frame server's wire protocol (gsr_b200/frame_server.py vs gaussian_renderer/network_gui.py:26-85) over a loopback
socket with a scripted client.  CPU: the frame payload is synthetic bytes (the device conversion has its own tests)."""
import json
import socket
import threading

import numpy as np

from gsr_b200 import frame_server


def _request(w, h, train=True, keep_alive=True):
    view = np.eye(4, dtype=np.float32)
    view[3, :3] = [0.1, -0.2, 3.0]
    return dict(resolution_x=w, resolution_y=h, train=int(train), fov_y=0.6, fov_x=0.8, z_near=0.01, z_far=100.0,
                shs_python=0, rot_scale_python=1, keep_alive=int(keep_alive), scaling_modifier=0.75,
                view_matrix=view.ravel().tolist(), view_projection_matrix=(view * 2).ravel().tolist())


def _client(port, requests, frame_len, results):
    s = socket.create_connection(("127.0.0.1", port), timeout=10)
    try:
        for req in requests:
            body = json.dumps(req).encode("utf-8")
            # split the message to exercise short reads on the server side
            s.sendall(len(body).to_bytes(4, "little") + body[:7])
            s.sendall(body[7:])
            want = frame_len if req["resolution_x"] and req["resolution_y"] else 0
            frame = b""
            while len(frame) < want:
                frame += s.recv(want - len(frame))
            n = int.from_bytes(s.recv(4), "little")
            verify = b""
            while len(verify) < n:
                verify += s.recv(n - len(verify))
            results.append((frame, verify.decode("ascii")))
    finally:
        s.close()


def test_protocol_roundtrip_over_loopback():
    srv = frame_server.FrameServer()
    port = srv.init("127.0.0.1", 0)
    assert not srv.try_connect()            # nobody there yet: non-blocking, like network_gui.try_connect
    W, H = 8, 6
    frame = bytes(range(W * H * 3))
    results = []
    reqs = [_request(W, H, train=False), _request(0, 0, train=False), _request(W, H, train=True)]
    th = threading.Thread(target=_client, args=(port, reqs, len(frame), results))
    th.start()
    import time
    for _ in range(200):
        if srv.try_connect():
            break
        time.sleep(0.01)
    assert srv.conn is not None
    seen = []

    # serve_pending with a stub renderer: answers until a request allows training to continue (train.py:74-87)
    import gsr_b200.frames as frames
    orig = frames.frame_bytes
    frames.frame_bytes = lambda img: memoryview(frame)
    try:
        sent = srv.serve_pending(lambda cam, mod, shs_py, rs_py: seen.append((cam, mod, shs_py, rs_py)) or "image", "/data/scene")
    finally:
        frames.frame_bytes = orig
    th.join(timeout=10)
    srv.close()
    assert sent == 2 and len(results) == 3
    assert results[0] == (frame, "/data/scene") and results[1] == (b"", "/data/scene") and results[2] == (frame, "/data/scene")
    cam, mod, shs_py, rs_py = seen[0]
    assert (cam.image_width, cam.image_height) == (W, H) and mod == 0.75 and shs_py is False and rs_py is True
    # network_gui.py:70-74: columns 1 and 2 of the view matrix and column 1 of the projection are negated
    assert np.allclose(cam.world_view_transform[3], [0.1, 0.2, -3.0, 1.0])
    assert np.allclose(cam.full_proj_transform[3], [0.2, 0.4, 6.0, 2.0])


def test_broken_connection_is_dropped():
    srv = frame_server.FrameServer()
    port = srv.init("127.0.0.1", 0)
    c = socket.create_connection(("127.0.0.1", port), timeout=10)
    import time
    for _ in range(200):
        if srv.try_connect():
            break
        time.sleep(0.01)
    c.sendall((100).to_bytes(4, "little") + b"{")       # announces 100 bytes, sends 1, hangs up
    c.close()
    assert srv.serve_pending(lambda *a: None, "x") == 0
    assert srv.conn is None                             # like the reference: network_gui.conn = None on any error
    srv.close()
