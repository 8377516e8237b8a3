"""Drop-in for the `rlgr` pybind11 module of python/PyRLGR (bindings.cpp:34-57).

class membuf: membuf() / membuf(buf), rlgrWrite(seq, flagSigned) -> ns, close(), get_buffer() ->
list[int], rlgrRead(N, flagSigned) -> (ns, list[int]), buffer_size().  The coding runs on the GPU
(raht_rlgr_encode / raht_rlgr_decode) and produces / consumes the reference bitstream byte for byte.
This list-based interface exists for parity with encode_3dgs.py:229-245; the fast path is
raht_b200.ops.rlgr_encode / rlgr_decode on device tensors (no Python lists).
grRead / grWrite (raw Golomb-Rice calls, unused by the drivers) are not provided.
"""
import time

import numpy as np
import torch

from raht_b200 import ops


class membuf:
    def __init__(self, buf=None):
        self._buf = bytes(bytearray(buf)) if buf is not None else b""
        self._streams = None

    def rlgrWrite(self, seq, flagSigned):
        a = np.asarray(seq, dtype=np.int64)
        if a.size and (a.min() < -2**31 or a.max() >= 2**31):
            raise OverflowError("rlgrWrite: values must fit int32")
        torch.cuda.synchronize()
        t0 = time.perf_counter_ns()
        q = torch.from_numpy(a.astype(np.int32)).cuda().reshape(1, -1)
        streams = ops.rlgr_encode(q, int(flagSigned))
        self._buf += streams.to_bytes()[0]
        return time.perf_counter_ns() - t0

    def close(self):
        pass          # streams are closed (zero-padded to a byte) by the encoder kernel

    def get_buffer(self):
        return list(self._buf)

    def buffer_size(self):
        return len(self._buf)

    def rlgrRead(self, N, flagSigned):
        t0 = time.perf_counter_ns()
        nb = len(self._buf)
        slot = max((nb + 8 + 7) // 8 * 8, 8)
        host = np.zeros(slot, dtype=np.uint8)
        host[:nb] = np.frombuffer(self._buf, dtype=np.uint8)
        buf = torch.from_numpy(host).cuda()
        nbytes = torch.tensor([nb], dtype=torch.int64, device=buf.device)
        q = ops.rlgr_decode(ops.RlgrStreams(buf, slot, nbytes, int(N), 1), int(flagSigned))
        out = q[0].cpu().tolist()
        return time.perf_counter_ns() - t0, out
