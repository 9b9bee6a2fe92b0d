"""The deployed inference loop, batched over many drones.

Reference: ``samples/inference_udp_FPGA_custom_dispatch/main.cpp`` -- one UDP socket (port 10001, :33), one client, one
frame per packet, the LSTM state carried from frame to frame (:217-221), the raw command post-processed into a velocity
(:381-417).  Wire format (:36-48, :320-370):

* request, 5424 bytes: 60x90 ``uint8`` depth image (5400) | desired velocity ``f32`` big-endian | position x ``f32``
  big-endian | quaternion 4 x ``f32`` big-endian;
* reply, 12 bytes: 3 x ``f32`` velocity command in the HOST's byte order (``htonf_noswap``, :357-369: no swap).

What is different here is the serving model, not the protocol: a B200 runs ~10^6 frames/s of this network, a drone sends
tens of frames per second, so the loop drains every datagram that is waiting, groups them into ONE batch (one frame per
drone per batch; a second frame of the same drone waits for the next batch, because it needs the first one's LSTM state),
keeps one (h, c) state per client address, and answers each client from the batched result.

Two reference details are kept behind flags because a drop-in may have to match them:

* ``prescale_velocity`` (default on): the loop divides the desired velocity by 10 before the model call (:180) although the
  model's ``forward`` divides by 10 again (``QAT/model.py:127``);
* ``quat_stride``: ``unpack_frame`` advances the quaternion read offset by ``sizeof(QUAT_SIZE / 4)`` = 8 instead of 4
  (:345), i.e. it reads components 0, 2 and two words beyond the packet.  The default here is the documented layout
  (stride 4); ``quat_stride=8`` reproduces the reference's reads (bytes beyond the packet read as 0).
"""
from __future__ import annotations

import select
import socket
import time
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

PORT = 10001                                   # main.cpp:33
IMAGE_H, IMAGE_W = 60, 90                      # tensor shape {1, 1, 60, 90} (main.cpp:174)
IMAGE_BYTES = IMAGE_H * IMAGE_W                # 5400
REQUEST_BYTES = IMAGE_BYTES + 4 + 4 + 16       # 5424
REPLY_BYTES = 12
MIN_X_VELOCITY_CMD = 1.0                       # main.cpp:404-405
CONTROL_THRESHOLD = 2.0


# ------------------------------------------------------------------------------------------------------------ wire format
def unpack_frames(packets: np.ndarray, quat_stride: int = 4):
    """[N, >=5424] uint8 datagrams -> (image u8 [N,60,90], desired velocity f32 [N], position x f32 [N], quaternion f32 [N,4])."""
    p = np.ascontiguousarray(packets, dtype=np.uint8)
    if p.ndim != 2 or p.shape[1] < REQUEST_BYTES:
        raise ValueError(f"expected [N, >= {REQUEST_BYTES}] uint8 packets, got {p.shape}")
    n = p.shape[0]
    img = p[:, :IMAGE_BYTES].reshape(n, IMAGE_H, IMAGE_W)
    be = np.dtype(">f4")
    vel = p[:, IMAGE_BYTES:IMAGE_BYTES + 4].copy().view(be).reshape(n).astype(np.float32)
    posx = p[:, IMAGE_BYTES + 4:IMAGE_BYTES + 8].copy().view(be).reshape(n).astype(np.float32)
    q0 = IMAGE_BYTES + 8
    if quat_stride == 4:
        quat = p[:, q0:q0 + 16].copy().view(be).reshape(n, 4).astype(np.float32)
    else:
        ext = np.zeros((n, q0 + 3 * quat_stride + 4), np.uint8)
        w = min(ext.shape[1], p.shape[1])
        ext[:, :w] = p[:, :w]
        quat = np.stack([ext[:, q0 + i * quat_stride:q0 + i * quat_stride + 4].copy().view(be).reshape(n) for i in range(4)],
                        axis=1).astype(np.float32)
    return img, vel, posx, quat


def pack_request(image_u8: np.ndarray, desired_velocity: float, position_x: float, quaternion: Sequence[float]) -> bytes:
    """The client side of the protocol (what the simulator / drone sends): used by the tests and the load generator."""
    img = np.ascontiguousarray(image_u8, dtype=np.uint8).reshape(IMAGE_BYTES)
    tail = np.array([desired_velocity, position_x, *quaternion], dtype=">f4")
    return img.tobytes() + tail.tobytes()


def pack_replies(cmd: np.ndarray) -> List[bytes]:
    """[N,3] f32 -> N 12-byte replies, floats in host byte order (main.cpp:357-369 copies the bytes unswapped)."""
    c = np.ascontiguousarray(cmd, dtype=np.float32).reshape(-1, 3)
    return [c[i].tobytes() for i in range(c.shape[0])]


def final_velocity(raw: np.ndarray, desired_vel: np.ndarray, pos_x: np.ndarray) -> np.ndarray:
    """``calculate_final_velocity`` (main.cpp:381-417), vectorised over the batch, in float32 like the reference:
    clip x to [-1, 1]; normalise by the Euclidean norm when it is > 0; scale by the desired velocity; while
    ``pos_x < 2`` the x command is ``max(1, pos_x / 2 * desired_vel)``."""
    v = np.array(raw, dtype=np.float32).reshape(-1, 3).copy()
    dv = np.asarray(desired_vel, dtype=np.float32).reshape(-1)
    px = np.asarray(pos_x, dtype=np.float32).reshape(-1)
    v[:, 0] = np.clip(v[:, 0], np.float32(-1.0), np.float32(1.0))
    norm = np.sqrt(v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1] + v[:, 2] * v[:, 2], dtype=np.float32)
    nz = norm > 0
    v[nz] = v[nz] / norm[nz, None]
    v *= dv[:, None]
    ramp = px < np.float32(CONTROL_THRESHOLD)
    v[ramp, 0] = np.maximum(np.float32(MIN_X_VELOCITY_CMD), (px[ramp] / np.float32(CONTROL_THRESHOLD)) * dv[ramp])
    return v


# ------------------------------------------------------------------------------------------------------ per-client state
class ClientStates:
    """LSTM state of every known client in ONE pair of device tensors ``[L, slots, Hd]`` plus an ``addr -> slot`` table, so
    that a batch gathers its states with one ``index_select`` and stores them with one ``index_copy_`` (separate small tensors
    per client cost two ``torch.stack`` over the batch and two clones per client: thousands of tiny launches at a 4096-frame
    batch, several times the model step).  Reads like a mapping ``addr -> (h [L,Hd], c [L,Hd], last_seen)``."""

    def __init__(self, layers: int, hidden: int, device: torch.device, capacity: int = 64):
        self.L, self.Hd, self.device = int(layers), int(hidden), torch.device(device)
        self.h = torch.zeros((self.L, capacity, self.Hd), device=self.device)
        self.c = torch.zeros_like(self.h)
        self.slot: Dict[Tuple, int] = {}
        self.seen: Dict[Tuple, float] = {}
        self.free: List[int] = list(range(capacity - 1, -1, -1))

    def __len__(self) -> int:
        return len(self.slot)

    def __contains__(self, addr) -> bool:
        return addr in self.slot

    def __getitem__(self, addr):
        i = self.slot[addr]
        return self.h[:, i], self.c[:, i], self.seen[addr]

    def _grow(self, need: int) -> None:
        cap = self.h.shape[1]
        new = max(2 * cap, cap + need)
        for name in ("h", "c"):
            t = torch.zeros((self.L, new, self.Hd), device=self.device)
            t[:, :cap] = getattr(self, name)
            setattr(self, name, t)
        self.free = list(range(new - 1, cap - 1, -1)) + self.free

    def slots_for(self, addrs: Sequence[Tuple], now: float) -> torch.Tensor:
        """Slot index of every address (device int64 [n]); unknown clients get a zeroed slot."""
        fresh = [a for a in addrs if a not in self.slot]
        if len(fresh) > len(self.free):
            self._grow(len(fresh) - len(self.free))
        new_slots = []
        for a in fresh:
            self.slot[a] = self.free.pop()
            new_slots.append(self.slot[a])
        if new_slots:                                   # a recycled slot still holds its previous owner's state
            z = torch.tensor(new_slots, device=self.device)
            self.h.index_fill_(1, z, 0.0)
            self.c.index_fill_(1, z, 0.0)
        for a in addrs:
            self.seen[a] = now
        return torch.tensor([self.slot[a] for a in addrs], dtype=torch.int64, device=self.device)

    def gather(self, idx: torch.Tensor):
        return self.h.index_select(1, idx), self.c.index_select(1, idx)

    def scatter(self, idx: torch.Tensor, h: torch.Tensor, c: torch.Tensor) -> None:
        self.h.index_copy_(1, idx, h.to(self.h.dtype))
        self.c.index_copy_(1, idx, c.to(self.c.dtype))

    def expire(self, now: float, ttl_s: float) -> None:
        for a in [a for a, t in self.seen.items() if now - t > ttl_s]:
            self.free.append(self.slot.pop(a))
            del self.seen[a]


# ------------------------------------------------------------------------------------------------------------ the server
class DroneServer:
    """Batched UDP inference service.

    ``model([img f32 [B,1,60,90], vel f32 [B,1], quat f32 [B,4], (h, c) or None]) -> (cmd [B,3], (h [L,B,Hd], c [L,B,Hd]))`` is
    ``ITAViTLSTM`` (or anything with that signature); tensors live on ``device``.
    """

    def __init__(self, model: Callable, device: torch.device, host: str = "0.0.0.0", port: int = PORT, max_batch: int = 4096,
                 prescale_velocity: bool = True, quat_stride: int = 4, lstm_layers: int = 3, lstm_hidden: int = 128,
                 state_ttl_s: float = 60.0):
        self.model, self.device = model, torch.device(device)
        self.max_batch, self.prescale_velocity, self.quat_stride = int(max_batch), bool(prescale_velocity), int(quat_stride)
        self.L, self.Hd, self.state_ttl_s = int(lstm_layers), int(lstm_hidden), float(state_ttl_s)
        self.state = ClientStates(self.L, self.Hd, self.device)                    # addr -> (h [L,Hd], c [L,Hd], last seen)
        self.sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        self.sock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        try:
            self.sock.setsockopt(socket.SOL_SOCKET, socket.SO_RCVBUF, 64 << 20)
        except OSError:
            pass
        self.sock.bind((host, port))
        self.sock.setblocking(False)
        self.address = self.sock.getsockname()
        self.frames = 0
        self.batches = 0
        pin = self.device.type == "cuda"
        self._packets = torch.empty((self.max_batch, REQUEST_BYTES), dtype=torch.uint8, pin_memory=pin)

    # -------------------------------------------------------------------------------------------------------- one batch
    def infer(self, packets: np.ndarray, addrs: Sequence[Tuple]) -> List[bytes]:
        """One frame per distinct client: unpack, run the model once on the whole batch with each client's LSTM state, store
        the new states, return the 12-byte replies in packet order."""
        n = len(addrs)
        if len(set(addrs)) != n:
            raise ValueError("a batch holds at most one frame per client (the second one needs the first one's state)")
        img_u8, vel, posx, quat = unpack_frames(packets, self.quat_stride)
        dev = self.device
        # the image goes over PCIe as bytes (5400 B per frame) and is scaled to [0, 1] on the device (main.cpp:168)
        img = torch.from_numpy(np.ascontiguousarray(img_u8)).to(dev, non_blocking=True).float().div_(255.0).unsqueeze(1)
        v_in = vel / np.float32(10.0) if self.prescale_velocity else vel
        v_t = torch.from_numpy(np.ascontiguousarray(v_in.reshape(n, 1))).to(dev, non_blocking=True)
        q_t = torch.from_numpy(np.ascontiguousarray(quat)).to(dev, non_blocking=True)
        idx = self.state.slots_for(addrs, time.monotonic())        # new clients start from a zero state
        h, c = self.state.gather(idx)                              # one index_select each, whatever the batch
        with torch.no_grad():
            cmd, (hn, cn) = self.model([img, v_t, q_t, (h, c)])
        self.state.scatter(idx, hn, cn)                            # one index_copy_ each
        out = final_velocity(cmd.float().cpu().numpy(), vel, posx)
        self.frames += n
        self.batches += 1
        return pack_replies(out)

    def _drain(self, timeout_s: float):
        """Every datagram that is waiting (up to ``max_batch``), oldest first; malformed ones are dropped."""
        got: List[Tuple[bytes, Tuple]] = []
        r, _, _ = select.select([self.sock], [], [], timeout_s)
        if not r:
            return got
        while len(got) < self.max_batch:
            try:
                data, addr = self.sock.recvfrom(8192)          # the reference's buffer size (main.cpp:152)
            except (BlockingIOError, InterruptedError):
                break
            if len(data) >= REQUEST_BYTES:
                got.append((data, addr))
        return got

    def poll(self, timeout_s: float = 0.05) -> int:
        """Receive what is waiting, answer it; returns the number of frames served.  Frames of the same client are served in
        arrival order, one per batch."""
        pending = self._drain(timeout_s)
        served = 0
        while pending:
            seen, batch, rest = set(), [], []
            for item in pending:
                (rest if item[1] in seen else batch).append(item)
                seen.add(item[1])
            buf = self._packets[:len(batch)].numpy()
            for i, (data, _) in enumerate(batch):
                buf[i] = np.frombuffer(data, dtype=np.uint8, count=REQUEST_BYTES)
            replies = self.infer(buf, [a for _, a in batch])
            for rep, (_, addr) in zip(replies, batch):
                try:
                    self.sock.sendto(rep, addr)
                except OSError:
                    pass
            served += len(batch)
            pending = rest
        return served

    def expire(self) -> None:
        self.state.expire(time.monotonic(), self.state_ttl_s)

    def serve(self, max_frames: Optional[int] = None, idle_timeout_s: Optional[float] = None) -> int:
        """The loop (``while (true)`` at main.cpp:150); ends after ``max_frames`` frames or ``idle_timeout_s`` without traffic
        when given."""
        last = time.monotonic()
        start_frames = self.frames
        while max_frames is None or self.frames - start_frames < max_frames:
            if self.poll(0.05):
                last = time.monotonic()
            elif idle_timeout_s is not None and time.monotonic() - last > idle_timeout_s:
                break
            if self.batches % 1024 == 0:
                self.expire()
        return self.frames - start_frames

    def close(self) -> None:
        self.sock.close()


def main(argv: Optional[Sequence[str]] = None) -> int:
    import argparse
    ap = argparse.ArgumentParser(description="ITAViTLSTM UDP inference service (protocol of samples/inference_udp_FPGA_custom_dispatch/main.cpp)")
    ap.add_argument("--checkpoint", required=True, help="converted int8 state dict (torch.save) with the float tensors")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=PORT)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--max-batch", type=int, default=4096)
    ap.add_argument("--quat-stride", type=int, default=4, choices=(4, 8))
    args = ap.parse_args(argv)
    from .model import ITAViTLSTM
    dev = torch.device("cuda", args.device)
    try:                                      # plain tensors / packed-parameter tuples load without running pickled code
        sd = dict(torch.load(args.checkpoint, weights_only=True))
    except Exception as e:
        raise SystemExit(f"{args.checkpoint}: not loadable with weights_only=True ({type(e).__name__}: {e}); re-save the converted "
                         "model's state_dict() from a trusted environment, this loader does not unpickle arbitrary objects")
    model = ITAViTLSTM.from_state_dict(sd).to(dev)
    srv = DroneServer(model, dev, args.host, args.port, args.max_batch, quat_stride=args.quat_stride)
    print(f"serving on {srv.address[0]}:{srv.address[1]} (device {dev})", flush=True)
    try:
        srv.serve()
    except KeyboardInterrupt:
        pass
    print(f"{srv.frames} frames in {srv.batches} batches")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
