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
