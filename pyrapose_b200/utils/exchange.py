"""Peer-to-peer gradient exchange of the data-parallel training step (C-ABI: ppb_p2p_*; csrc/exchange.cu).

The reference trains data-parallel with keras.utils.multi_gpu_model (PyraPose/bin/train.py:82-86): replica outputs are
concatenated and one gradient is applied. Here every replica is a process with its own GPU and
the replica gradients have to be summed. The default is NCCL's all-reduce, bucket by bucket, overlapped with the
backward pass; this module is the alternative that uses no SM while the backward pass runs (compile(exchange='p2p');
measured equal to NCCL at 2 GPUs and slower at 8, see csrc/exchange.cu):

  * every replica owns a staging buffer [2 parities][world][n] float32 and maps its peers' buffers (CUDA IPC handles
    exchanged once over torch.distributed);
  * push(): as soon as a gradient bucket is complete, it is copied into slot [parity][my rank] of EVERY peer's staging
    buffer with asynchronous device-to-device copies on side streams (copy engines over NVLink / NVSwitch);
  * finish(): a one-element all-reduce ordered behind the copies is the cross-replica "all pushes have landed" signal,
    then ONE kernel adds the world slices in rank order into the gradient buffer (identical on every replica).

Two parities alternate by step: a peer can only overwrite slot set p again after the signal of the step in between,
which this replica joins only after it has summed set p.

Single node only (peer mappings); `available()` says whether the replicas of a process group can use it."""
import ctypes

import numpy as np

from .. import _native as nat


def available(group=None):
    """True when torch.distributed is initialised with more than one rank, all on this node, each on its own GPU."""
    import os
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return False
    world = dist.get_world_size(group)
    if world < 2 or world > 16 or not torch.cuda.is_available():
        return False
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))
    return local_world == world and torch.cuda.device_count() >= world


class P2PExchange(object):
    def __init__(self, n, device, group=None, n_copy_streams=4):
        import torch
        import torch.distributed as dist
        self.lib = nat.load()
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.n = int(n)
        self.device = torch.device(device)
        self.slot_bytes = ((self.n * 4 + 255) // 256) * 256   # 256-byte slots keep every slice 16-byte aligned
        self.set_bytes = self.slot_bytes * self.world
        self._local = ctypes.c_void_p()
        self._peers = {}
        self.step = 0
        self.ok, self.error = True, None
        with torch.cuda.device(self.device):
            # local part; a failure here must not keep this replica out of the collective below
            mine = None
            try:
                nat.check(self.lib.ppb_p2p_alloc(ctypes.c_size_t(2 * self.set_bytes), ctypes.byref(self._local)))
                handle = (ctypes.c_ubyte * 64)()
                nat.check(self.lib.ppb_p2p_export(self._local, handle))
                mine = (bytes(bytearray(handle)), int(self.device.index))
            except Exception as e:
                self.ok, self.error = False, e
            everyone = [None] * self.world
            dist.all_gather_object(everyone, mine, group=group)
            if any(x is None for x in everyone):
                self.ok = False
                self.error = self.error or RuntimeError("a peer could not export its staging buffer")
            elif len(set(d for _, d in everyone)) != self.world:
                self.ok = False
                self.error = RuntimeError("one GPU per replica on one node is required, got devices %s"
                                          % ([d for _, d in everyone],))
            else:
                try:
                    for r, (h, _) in enumerate(everyone):
                        if r == self.rank:
                            continue
                        p = ctypes.c_void_p()
                        buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                        nat.check(self.lib.ppb_p2p_import(buf, ctypes.byref(p)))
                        self._peers[r] = p
                except Exception as e:
                    self.ok, self.error = False, e
            self.streams = [torch.cuda.Stream(device=self.device) for _ in range(max(1, int(n_copy_streams)))]
            self._flag = torch.zeros(1, dtype=torch.float32, device=self.device)
        # The caller now agrees with the other replicas (all-reduce of `ok`) whether everyone is connected; that
        # collective is also the barrier after which pushes may start.

    def _slot(self, base, parity, src_rank):
        return base + parity * self.set_bytes + src_rank * self.slot_bytes

    def push(self, grads, off, cnt):
        """Copies grads[off:off+cnt] (complete on the CURRENT stream at this point) into this replica's slot of every
        peer's staging buffer. Returns immediately."""
        import torch
        if cnt <= 0:
            return
        parity = self.step & 1
        main = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(main)
        src = grads.data_ptr() + 4 * int(off)
        for k, (r, base) in enumerate(sorted(self._peers.items())):
            st = self.streams[k % len(self.streams)]
            st.wait_event(ready)
            dst = self._slot(base.value, parity, self.rank) + 4 * int(off)
            nat.check(self.lib.ppb_p2p_copy(ctypes.c_void_p(dst), ctypes.c_void_p(src), ctypes.c_size_t(4 * int(cnt)),
                                            ctypes.c_void_p(st.cuda_stream)))

    def finish(self, grads):
        """After the last push of the step: waits (on the device) until every replica's pushes have landed, then
        grads = sum over replicas in rank order. Everything is stream-ordered; no host synchronisation."""
        import torch
        import torch.distributed as dist
        parity = self.step & 1
        main = torch.cuda.current_stream(self.device)
        s0 = self.streams[0]
        for st in self.streams[1:]:
            e = torch.cuda.Event()
            e.record(st)
            s0.wait_event(e)
        with torch.cuda.stream(s0):
            # ordered behind this replica's copies; completes only when every replica has joined, i.e. when every
            # replica's copies are done
            work = dist.all_reduce(self._flag, group=self.group, async_op=True)
        work.wait()   # the current (main) stream waits for it on the device
        srcs = (ctypes.c_void_p * self.world)()
        for r in range(self.world):
            srcs[r] = grads.data_ptr() if r == self.rank else self._slot(self._local.value, parity, r)
        nat.check(self.lib.ppb_p2p_sum_f32(ctypes.c_void_p(grads.data_ptr()), srcs, self.world,
                                           ctypes.c_longlong(self.n), ctypes.c_void_p(main.cuda_stream)))
        self.step += 1

    def close(self):
        import torch
        if getattr(self, "_closed", False):
            return
        self._closed = True
        try:
            with torch.cuda.device(self.device):
                torch.cuda.synchronize()
                for p in self._peers.values():
                    self.lib.ppb_p2p_release(p)
                self._peers = {}
                if self._local:
                    self.lib.ppb_p2p_free(self._local)
                    self._local = ctypes.c_void_p()
        except Exception:
            pass

    def __del__(self):
        self.close()
