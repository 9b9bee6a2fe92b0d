"""The UDP serving loop: wire format, velocity post-processing and the batching / per-client state logic.

Reference behaviour: samples/inference_udp_FPGA_custom_dispatch/main.cpp (:36-48 sizes, :320-355 unpack, :357-369 reply,
:381-417 post-processing, :150-226 the loop).  The CPU tests drive the server with a stub model over the loopback interface
(the stub is not a compute path of the product: it makes the state bookkeeping observable); the GPU test serves the real
``ITAViTLSTM``.
"""
import math
import os
import socket
import struct
import threading

import numpy as np
import pytest

from conftest import GOLDEN

torch = pytest.importorskip("torch")


def _serve_mod():
    import importlib
    return importlib.import_module("drone-oa-iree-vit-accelerator_b200.serve")


def _final_velocity_scalar(raw, desired, pos_x):
    """Independent scalar restatement of main.cpp:381-417 in float32."""
    f = np.float32
    v = [f(raw[0]), f(raw[1]), f(raw[2])]
    v[0] = min(max(v[0], f(-1.0)), f(1.0))
    norm = f(math.sqrt(f(f(v[0] * v[0]) + f(v[1] * v[1])) + f(v[2] * v[2])))
    if norm > 0:
        v = [f(x / norm) for x in v]
    v = [f(x * f(desired)) for x in v]
    if f(pos_x) < f(2.0):
        v[0] = max(f(1.0), f(f(f(pos_x) / f(2.0)) * f(desired)))
    return v


def test_wire_format_round_trip_and_byte_order():
    S = _serve_mod()
    assert S.REQUEST_BYTES == 5424 and S.REPLY_BYTES == 12 and S.PORT == 10001
    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, size=(5, 60, 90), dtype=np.uint8)
    vel = rng.uniform(1, 9, 5).astype(np.float32)
    px = rng.uniform(0, 60, 5).astype(np.float32)
    quat = rng.normal(size=(5, 4)).astype(np.float32)
    pk = [S.pack_request(imgs[i], vel[i], px[i], quat[i]) for i in range(5)]
    assert all(len(p) == 5424 for p in pk)
    # big-endian floats after the image, as the reference's unpack_frame expects
    assert struct.unpack(">f", pk[2][5400:5404])[0] == vel[2] and struct.unpack(">f", pk[2][5404:5408])[0] == px[2]
    arr = np.stack([np.frombuffer(p, np.uint8) for p in pk])
    i2, v2, p2, q2 = S.unpack_frames(arr)
    assert np.array_equal(i2, imgs) and np.array_equal(v2, vel) and np.array_equal(p2, px) and np.array_equal(q2, quat)
    # the reference's own read pattern (offset += 8 per component): components 0 and 2, then beyond the packet
    _, _, _, q8 = S.unpack_frames(arr, quat_stride=8)
    assert np.array_equal(q8[:, 0], quat[:, 0]) and np.array_equal(q8[:, 1], quat[:, 2]) and not q8[:, 2:].any()
    # replies: host byte order, no swap
    rep = S.pack_replies(np.array([[1.5, -2.25, 3.0]], np.float32))
    assert rep[0] == np.array([1.5, -2.25, 3.0], np.float32).tobytes() and len(rep[0]) == 12
    with pytest.raises(ValueError):
        S.unpack_frames(np.zeros((2, 100), np.uint8))


def test_final_velocity_matches_scalar_restatement():
    S = _serve_mod()
    rng = np.random.default_rng(1)
    raw = rng.normal(scale=1.5, size=(2000, 3)).astype(np.float32)
    raw[:5] = 0.0                                           # zero norm: no division
    raw[5:10, 0] = 7.0                                      # clipped
    dv = rng.uniform(0.5, 9, 2000).astype(np.float32)
    px = rng.uniform(-1, 6, 2000).astype(np.float32)
    got = S.final_velocity(raw, dv, px)
    want = np.array([_final_velocity_scalar(raw[i], dv[i], px[i]) for i in range(2000)], np.float32)
    assert np.allclose(got, want, rtol=0, atol=1e-6), np.abs(got - want).max()
    ramp = px < 2
    assert (got[ramp, 0] >= 1.0).all()


class _StubModel:
    """cmd = [mean(img), 10 * vel, h[0, :, 0]] ; h' = h + 1, c' = c + 2: makes the state bookkeeping visible in the replies."""

    def __call__(self, X):
        img, vel, quat, (h, c) = X
        cmd = torch.stack([img.mean(dim=(1, 2, 3)), vel[:, 0] * 10.0, h[0, :, 0]], dim=1)
        return cmd, (h + 1.0, c + 2.0)


def test_server_batches_clients_and_keeps_per_client_state():
    S = _serve_mod()
    srv = S.DroneServer(_StubModel(), torch.device("cpu"), host="127.0.0.1", port=0, max_batch=8, prescale_velocity=True)
    port = srv.address[1]
    clients = [socket.socket(socket.AF_INET, socket.SOCK_DGRAM) for _ in range(3)]
    for c in clients:
        c.bind(("127.0.0.1", 0))
        c.settimeout(5.0)
    img = np.full((60, 90), 51, np.uint8)                   # mean 0.2 after /255
    # burst: two frames from every client before the server looks at the socket
    for rnd in range(2):
        for k, c in enumerate(clients):
            c.sendto(S.pack_request(img, 4.0 + k, 10.0, [1, 0, 0, 0]), ("127.0.0.1", port))
    served = 0
    while served < 6:
        served += srv.poll(1.0)
    assert srv.frames == 6 and srv.batches >= 2             # same client twice => at least two batches
    assert len(srv.state) == 3
    for k, c in enumerate(clients):
        r1 = np.frombuffer(c.recvfrom(64)[0], np.float32)
        r2 = np.frombuffer(c.recvfrom(64)[0], np.float32)
        dv = np.float32(4.0 + k)
        # stub output [0.2, (dv/10)*10, state]; state 0 on the first frame, 1 on the second
        want1 = S.final_velocity(np.array([[0.2, dv, 0.0]], np.float32), [dv], [10.0])[0]
        want2 = S.final_velocity(np.array([[0.2, dv, 1.0]], np.float32), [dv], [10.0])[0]
        assert np.allclose(r1, want1, atol=1e-6) and np.allclose(r2, want2, atol=1e-6), (k, r1, want1, r2, want2)
        assert float(srv.state[c.getsockname()][0][0, 0]) == 2.0 and float(srv.state[c.getsockname()][1][0, 0]) == 4.0
    # one frame per client per batch is an invariant of infer()
    with pytest.raises(ValueError):
        srv.infer(np.zeros((2, 5424), np.uint8), [("a", 1), ("a", 1)])
    # short datagrams are dropped, the loop keeps running
    clients[0].sendto(b"junk", ("127.0.0.1", port))
    assert srv.poll(0.2) == 0
    srv.state.slots_for([("10.0.0.1", 1)], now=-1e9)       # a client last seen long ago
    assert ("10.0.0.1", 1) in srv.state and len(srv.state) == 4
    srv.expire()
    assert ("10.0.0.1", 1) not in srv.state and len(srv.state) == 3
    # the freed slot is recycled with a zero state, and the table grows past its first capacity
    many = [("10.0.1.%d" % i, 7) for i in range(200)]
    idx = srv.state.slots_for(many, now=0.0)
    assert len(set(idx.tolist())) == 200 and len(srv.state) == 203
    h0, c0 = srv.state.gather(idx)
    assert float(h0.abs().max()) == 0.0 and float(c0.abs().max()) == 0.0
    for c in clients:
        c.close()
    srv.close()


@pytest.mark.gpu
@pytest.mark.skipif(not torch.cuda.is_available(), reason="needs a CUDA device")
def test_server_serves_the_real_model_on_gpu():
    """Three drones, two frames each, over loopback: every reply equals the post-processed output of a direct
    ``ITAViTLSTM`` call with that drone's own recurrent state."""
    import importlib
    import sys
    sys.path.insert(0, GOLDEN)
    import seeded
    S = _serve_mod()
    M = importlib.import_module("drone-oa-iree-vit-accelerator_b200.model")
    dev = torch.device("cuda", 0)
    sd = dict(torch.load(os.path.join(GOLDEN, "qat_state_E64.pt"), weights_only=False))
    for name, shape in M.reference_float_shapes(1, 64).items():
        sd[name] = torch.from_numpy(seeded.float_param(name, shape))
    model = M.ITAViTLSTM.from_state_dict(sd).to(dev)
    srv = S.DroneServer(model, dev, host="127.0.0.1", port=0, max_batch=16)
    port = srv.address[1]
    rng = np.random.default_rng(3)
    clients = [socket.socket(socket.AF_INET, socket.SOCK_DGRAM) for _ in range(3)]
    for c in clients:
        c.bind(("127.0.0.1", 0))
        c.settimeout(10.0)
    frames = rng.integers(0, 256, size=(2, 3, 60, 90), dtype=np.uint8)
    quat = rng.normal(size=(3, 4)).astype(np.float32)
    dvel = np.array([3.0, 5.0, 7.0], np.float32)
    posx = np.array([0.5, 10.0, 30.0], np.float32)
    t = threading.Thread(target=srv.serve, kwargs={"max_frames": 6, "idle_timeout_s": 20.0})
    t.start()
    hidden = None
    for step in range(2):
        for k, c in enumerate(clients):
            c.sendto(S.pack_request(frames[step, k], dvel[k], posx[k], quat[k]), ("127.0.0.1", port))
        got = np.stack([np.frombuffer(c.recvfrom(64)[0], np.float32) for c in clients])
        img = torch.from_numpy(frames[step]).to(dev).float().div(255.0).unsqueeze(1)
        cmd, hidden = model([img, torch.from_numpy(dvel / np.float32(10.0)).to(dev).reshape(3, 1), torch.from_numpy(quat).to(dev), hidden])
        want = S.final_velocity(cmd.cpu().numpy(), dvel, posx)
        assert np.allclose(got, want, rtol=0, atol=5e-3), (step, np.abs(got - want).max())   # a different batch split may flip an LSB of a quantised token
    t.join(30.0)
    assert not t.is_alive() and srv.frames == 6
    for c in clients:
        c.close()
    srv.close()
