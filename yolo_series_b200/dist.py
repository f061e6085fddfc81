"""Image-sharded multi-GPU inference: one process per GPU, one replica per GPU,
contiguous split of the batch along dim 0, and exactly ONE collective per step -- a gather of the
fixed-size detection buffers ``(B_local, max_det, 6) fp32 + counts`` to every rank (NCCL over
NVLink; gloo on CPU for the host-logic tests).  The reference has no multi-GPU inference at all
(utils/torch_utils.py:86 returns cuda:0 only); images are independent units at every stage of the
path, so no data-path collective exists inside the forward / decode / NMS.
"""
import torch
import torch.distributed as dist


def shard_bounds(B, rank, world):
    """Contiguous [lo, hi) slice of a batch of B images owned by ``rank`` (remainder to low ranks)."""
    base, rem = divmod(B, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_detections(out, counts):
    """(B, max_det, 6) fp32 + (B,) int32 -> one (B, max_det*6 + 1) fp32 message (count bit-cast)."""
    B = out.shape[0]
    msg = torch.empty((B, out.shape[1] * 6 + 1), dtype=torch.float32, device=out.device)
    msg[:, :-1] = out.reshape(B, -1)
    msg[:, -1] = counts.to(torch.int32).view(torch.float32)
    return msg


def unpack_detections(msg, max_det):
    counts = msg[:, -1].contiguous().view(torch.int32)
    return msg[:, :-1].reshape(msg.shape[0], max_det, 6), counts


def gather_detections(out, counts, group=None):
    """All ranks contribute equally sized shards; returns (out_all, counts_all) in global image order."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return out, counts
    world = dist.get_world_size(group)
    msg = pack_detections(out, counts)
    allmsg = torch.empty((world * msg.shape[0], msg.shape[1]), dtype=msg.dtype, device=msg.device)
    dist.all_gather_into_tensor(allmsg, msg, group=group)
    return unpack_detections(allmsg, out.shape[1])


def to_list(out, counts):
    c = counts.tolist()
    return [out[i, :c[i]] for i in range(out.shape[0])]


class PeerGather:
    """The gather of SURVEY.md 8(e) fused into the NMS kernel (``yb_decode_nms_gather``): every rank owns one
    cudaMalloc'd region, exported to the other ranks of the box through CUDA IPC, holding two buffer sets of
    ``(world, B, max_det, 6)`` fp32 detections + ``(world, B)`` int32 counts + ``(world,)`` arrival flags.  The NMS
    kernel of rank r stores each kept row into block r of EVERY rank's current set over NVLink and raises flag r there;
    ``finish`` polls the local flags (``yb_peer_wait``) and copies the set out.  Protocol for slot reuse: set
    ``step & 1`` is written at step s and again at s + 2; a rank can only start step s + 2 after its own ``finish(s + 1)``,
    which needs every peer's flag for s + 1, which each peer raises after its own ``finish(s)`` copy -- all on the
    compute stream, so nobody overwrites a set that is still being read.
    """

    def __init__(self, B, max_det, group=None, device=None, timeout_s=None):
        """``timeout_s`` (default: env YB_PEER_TIMEOUT_S or 20): how long ``finish`` waits for a peer before it poisons that
        peer's block (counts = -1) and sets the status word; ``Model.detect`` / ``detect_batches`` raise on a poisoned
        block wherever they already read the counts back."""
        import ctypes as C
        import os

        from . import _lib
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError('PeerGather needs an initialised torch.distributed process group')
        self.group = group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > _lib.MAX_PEERS:
            raise ValueError(f'PeerGather supports up to {_lib.MAX_PEERS} GPUs of one box')
        self.B, self.max_det = int(B), int(max_det)
        self.timeout_ms = int(1000 * float(timeout_s if timeout_s is not None else os.environ.get('YB_PEER_TIMEOUT_S', 20)))
        self.device = device if device is not None else torch.device('cuda', torch.cuda.current_device())
        self.lib = _lib.lib()
        w = self.world
        self.out_floats = w * self.B * self.max_det * 6
        self.cnt_ints = w * self.B
        # region layout (bytes): [set0 out | set1 out | set0 counts | set1 counts | set0 flags | set1 flags | done counter]
        al = lambda n: (n + 255) // 256 * 256
        self.off_out = [0, al(self.out_floats * 4)]
        base = 2 * al(self.out_floats * 4)
        self.off_cnt = [base, base + al(self.cnt_ints * 4)]
        base += 2 * al(self.cnt_ints * 4)
        self.off_flag = [base, base + 256]
        self.off_done = base + 512
        self.bytes = base + 768
        # Every rank must reach the same verdict (a rank that failed alone would leave the others in a barrier), so the
        # outcome of each stage is agreed on through all_gather_object before anyone proceeds.
        ptr, handle, err = C.c_void_p(), C.create_string_buffer(64), None
        try:
            with torch.cuda.device(self.device):
                _lib.check(self.lib.yb_peer_alloc(self.bytes, C.byref(ptr), handle), 'yb_peer_alloc')
        except RuntimeError as ex:
            err = str(ex)
        self.local = ptr.value
        handles = [None] * w
        dist.all_gather_object(handles, (handle.raw, err), group=group)
        self.peers = []
        if all(h[1] is None for h in handles):
            for r in range(w):
                if r == self.rank:
                    self.peers.append(self.local)
                    continue
                p = C.c_void_p()
                try:
                    with torch.cuda.device(self.device):
                        _lib.check(self.lib.yb_peer_open(handles[r][0], C.byref(p)), f'yb_peer_open(rank {r})')
                except RuntimeError as ex:
                    err = str(ex)
                    break
                self.peers.append(p.value)
        verdicts = [None] * w
        dist.all_gather_object(verdicts, err or next((h[1] for h in handles if h[1]), None), group=group)
        bad = next((v for v in verdicts if v), None)
        if bad:                                # same decision on every rank
            for r, p in enumerate(self.peers):
                if r != self.rank:
                    self.lib.yb_peer_close(C.c_void_p(p))
            if self.local:
                self.lib.yb_peer_free(C.c_void_p(self.local))
            self.peers = None
            raise RuntimeError(f'PeerGather: peer memory is not available on this box ({bad}); use dist.gather_detections')
        self.step = 0
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._views = [self._view_set(s) for s in (0, 1)]

    def _wrap(self, addr, shape, typestr, dtype):
        class _Cai:
            pass
        o = _Cai()
        o.__cuda_array_interface__ = {'shape': tuple(shape), 'typestr': typestr, 'data': (int(addr), False), 'version': 2}
        t = torch.as_tensor(o, device=self.device)
        assert t.dtype == dtype and t.data_ptr() == addr
        return t

    def _view_set(self, s):
        w = self.world
        out = self._wrap(self.local + self.off_out[s], (w * self.B, self.max_det, 6), '<f4', torch.float32)
        cnt = self._wrap(self.local + self.off_cnt[s], (w * self.B,), '<i4', torch.int32)
        return out, cnt

    def begin(self, B, max_det):
        """The table for the next NMS launch: (``_lib.PeerGatherC``, slot, step)."""
        import ctypes as C

        from . import _lib
        if (B, max_det) != (self.B, self.max_det):
            raise ValueError(f'PeerGather was built for B={self.B}, max_det={self.max_det}; got {B}, {max_det}')
        self.step += 1
        s = self.step & 1
        t = _lib.PeerGatherC()
        t.n_peers, t.rank, t.step = self.world, self.rank, self.step
        for r, base in enumerate(self.peers):
            t.out[r] = base + self.off_out[s]
            t.counts[r] = base + self.off_cnt[s]
            t.flags[r] = base + self.off_flag[s]
        t.done_counter = self.local + self.off_done
        return t, s, self.step

    def finish(self, slot, step):
        """Wait (on the current stream) until every rank's block of set ``slot`` has arrived, then return private copies
        ``(out_all (world*B, max_det, 6), counts_all (world*B,))`` in global image order."""
        import ctypes as C

        from . import _lib
        with torch.cuda.device(self.device):
            _lib.check(self.lib.yb_peer_wait(C.c_void_p(self.local + self.off_flag[slot]), self.world, step,
                                             C.c_void_p(self.status.data_ptr()),
                                             C.c_void_p(self.local + self.off_cnt[slot]), self.B, self.timeout_ms,
                                             _lib.stream_ptr()), 'yb_peer_wait')
        out, cnt = self._views[slot]
        return out.clone(), cnt.clone()

    @staticmethod
    def raise_if_poisoned(counts_host):
        """``counts_host``: the gathered per-image counts after a host read-back.  A block whose sender did not arrive in
        time carries -1 (k_peer_wait) -- never return it as detections."""
        if bool((counts_host < 0).any()):
            late = sorted({int(i) for i in (counts_host < 0).nonzero().flatten().tolist()})
            raise RuntimeError(f'PeerGather: timed out waiting for a peer rank (images {late[:4]}... carry no data); '
                               f'raise YB_PEER_TIMEOUT_S or check the slow rank')

    def check(self):
        """One 4-byte read-back: raises if a wait timed out (a peer never arrived)."""
        if int(self.status.item()) != 0:
            raise RuntimeError('PeerGather: timed out waiting for a peer rank')

    def close(self):
        import ctypes as C
        if getattr(self, 'peers', None) is None:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)      # nobody is still writing into a region that is about to go away
        for r, p in enumerate(self.peers):
            if r != self.rank:
                self.lib.yb_peer_close(C.c_void_p(p))
        self._views = None
        self.lib.yb_peer_free(C.c_void_p(self.local))
        self.peers = None
