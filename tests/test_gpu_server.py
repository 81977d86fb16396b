"""-m gpu: the drop-in boundary end to end — snapshot file -> C++ ngp::NerfRender facade -> render_server wire protocol
(/root/reference/src/render_server.cu:93-101: 64-byte row-major pose in, 3*W*H bytes u8 RGB out) -> same frame as the
C ABI called directly, and within the image tolerance of the CPU oracle."""
import os
import socket
import subprocess
import time

import numpy as np
import pytest

from nerf_cuda_b200 import NerfRender, synthetic as syn
from nerf_cuda_b200.snapshot import write_snapshot
from oracle import orc

pytestmark = pytest.mark.gpu
PKG = os.path.dirname(os.path.abspath(syn.__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _recv(sock, n):
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(n - len(buf))
        assert chunk, "server closed the connection early"
        buf += chunk
    return bytes(buf)


def test_snapshot_file_roundtrip_and_server(built_lib, tmp_path):
    sc = syn.make_scene("tiny", rgb_activation=syn.NRF_ACT_SIGMOID)
    path = str(tmp_path / "synthetic.msgpack")
    write_snapshot(path, sc.desc, sc.params, sc.grid, sc.aabb)
    W, H = 96, 64
    intr = syn.intrinsics(W, H)
    poses = [syn.look_at_pose(4.0, 20.0, az) for az in (15.0, 140.0, 260.0)]
    # direct: reload_network_from_file through the C ABI
    r = NerfRender(1)
    r.reload_network_from_file(path)
    r.set_resolution((W, H))
    direct = [r.render_frame(intr, p)[0].copy() for p in poses]
    r.close()
    O = orc.Oracle(sc.desc, sc.params, sc.grid, sc.aabb)
    ref = O.to_u8(O.render(intr, poses[0], W, H)[0]).reshape(H, W, 3)
    assert np.abs(direct[0].astype(int) - ref.astype(int)).max() <= 2
    # server: the reference's protocol; pose split across two TCP segments to exercise short reads
    port = _free_port()
    srv = subprocess.Popen([os.path.join(PKG, "render_server"), "--snapshot", path, "--port", str(port), "--width", str(W),
                            "--height", str(H), "--fl", str(float(intr[0])), "--cx", str(float(intr[2])), "--cy",
                            str(float(intr[3])), "--max-requests", str(len(poses))], stdout=subprocess.PIPE, text=True)
    try:
        line = srv.stdout.readline()
        assert "Awaiting connections" in line, line
        sock = socket.create_connection(("127.0.0.1", port), timeout=60)
        for k, p in enumerate(poses):
            raw = np.asarray(p, "<f4").tobytes()
            assert len(raw) == 64
            sock.sendall(raw[:20])
            time.sleep(0.02)
            sock.sendall(raw[20:])
            frame = np.frombuffer(_recv(sock, 3 * W * H), np.uint8).reshape(H, W, 3)
            assert np.array_equal(frame, direct[k])
        sock.close()
        assert srv.wait(timeout=30) == 0
    finally:
        if srv.poll() is None:
            srv.kill()


def test_reload_errors_are_the_references(built_lib, tmp_path):
    from nerf_cuda_b200 import NerfError
    r = NerfRender(1)
    with pytest.raises(NerfError) as e:
        r.reload_network_from_file(str(tmp_path / "nope.msgpack"))
    assert "does not exist" in str(e.value)
    with pytest.raises(NerfError) as e:
        r.set_resolution((8, 8))          # set_resolution before a model is loaded (nerf_render.cu:226)
    assert e.value.code == -6
    r.close()


def test_testbed_writes_ppm(built_lib, tmp_path):
    sc = syn.make_scene("tiny")
    path = str(tmp_path / "synthetic.msgpack")
    write_snapshot(path, sc.desc, sc.params, sc.grid, sc.aabb)
    out = str(tmp_path / "img")
    res = subprocess.run([os.path.join(PKG, "testbed"), path, "64", "48", out], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    assert "Process time" in res.stdout
    data = open(out + ".ppm", "rb").read()
    assert data.startswith(b"P6\n64 48\n255\n") and len(data) == len(b"P6\n64 48\n255\n") + 64 * 48 * 3


def test_server_batch_and_depth(built_lib, tmp_path):
    """render_server --batch N --depth: one request = N poses, the response = N x (rgb frame, depth frame), rendered
    through the pipelined nrf_render_batch.  Every frame must equal the one-at-a-time render_frame result byte for
    byte; --batch 0 is rejected instead of busy-looping."""
    sc = syn.make_scene("tiny16")
    path = str(tmp_path / "synthetic16.msgpack")
    write_snapshot(path, sc.desc, sc.params, sc.grid, sc.aabb)
    W, H, N = 80, 56, 4
    intr = syn.intrinsics(W, H)
    poses = [syn.look_at_pose(12.0, 20.0, az) for az in (10.0, 100.0, 190.0, 280.0, 300.0, 40.0, 77.0, 200.0)]
    r = NerfRender(1)
    r.reload_network_from_file(path)
    r.set_resolution((W, H))
    direct = [tuple(a.copy() for a in r.render_frame(intr, p)) for p in poses]
    rgb_b, dep_b = r.render_batch(intr, np.stack(poses[:N]), want_depth=True)
    for k in range(N):
        assert np.array_equal(rgb_b[k], direct[k][0]) and np.array_equal(dep_b[k], direct[k][1])
    r.close()
    assert subprocess.run([os.path.join(PKG, "render_server"), "--snapshot", path, "--batch", "0"],
                          capture_output=True, timeout=60).returncode == 2
    port = _free_port()
    srv = subprocess.Popen([os.path.join(PKG, "render_server"), "--snapshot", path, "--port", str(port), "--width", str(W),
                            "--height", str(H), "--fl", str(float(intr[0])), "--cx", str(float(intr[2])), "--cy",
                            str(float(intr[3])), "--batch", str(N), "--depth", "--max-requests", "2"],
                           stdout=subprocess.PIPE, text=True)
    try:
        assert "Awaiting connections" in srv.stdout.readline()
        sock = socket.create_connection(("127.0.0.1", port), timeout=60)
        for req in range(2):
            chunk = poses[req * N:(req + 1) * N]
            sock.sendall(np.asarray(np.stack(chunk), "<f4").tobytes())
            for k in range(N):
                frame = np.frombuffer(_recv(sock, 3 * W * H), np.uint8).reshape(H, W, 3)
                depth = np.frombuffer(_recv(sock, W * H), np.uint8).reshape(H, W)
                assert np.array_equal(frame, direct[req * N + k][0]) and np.array_equal(depth, direct[req * N + k][1])
        sock.close()
        assert srv.wait(timeout=30) == 0
    finally:
        if srv.poll() is None:
            srv.kill()
