"""The shuffle between ranks: one process per GPU, torch.distributed for the plumbing.

Replaces the reference's file-based all-to-all (every mapper writes <= R run files,
job.lua:208-221; every reducer lists and reads <= M of them, job.lua:253-260;
server.lua:279-329 discovers them).  Here every rank exports its combined
(key, count) records grouped by owner rank (owner = partition mod world_size, so each
reduce partition is owned by exactly one GPU), the segments move with ONE variable-size
all-to-all (NCCL over NVLink on GPUs; gloo in the CPU tests of the host logic), and the
owner merges what it received with the radix sort + segmented reduce.

Segment wire format (per source -> destination): n records of 16 bytes
{key word u64, count u64} followed by the bytes of the keys of 8+ bytes; a long key's
word is 1<<63 | offset-in-byte-block << 26 | length.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_splits(n_splits, world_size, rank):
    """Split ids of one rank: round-robin, like workers claiming map jobs in turn (task.lua:258-343)."""
    return list(range(rank, n_splits, world_size))


def owner_of_partition(part, world_size):
    return part % world_size


class _DeviceBytes:
    """Wraps a raw device pointer so torch can view it (no copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 2}


def device_bytes(ptr, n, device):
    if n == 0:
        return torch.empty(0, dtype=torch.uint8, device=device)
    return torch.as_tensor(_DeviceBytes(ptr, n), device=device)


def exchange_plan(send_bytes, send_nrec, group=None, device="cpu"):
    """Tells every rank what it will receive.  send_bytes / send_nrec: per destination.
    Returns (recv_bytes, recv_nrec, recv_off) as int64 numpy arrays (recv_off has W+1 entries)."""
    W = dist.get_world_size(group)
    meta = torch.tensor(np.stack([np.asarray(send_bytes, np.int64), np.asarray(send_nrec, np.int64)], 1),
                        dtype=torch.int64, device=device).contiguous()
    got = torch.empty_like(meta)
    dist.all_to_all_single(got, meta, group=group)
    got = got.cpu().numpy()
    recv_bytes, recv_nrec = got[:, 0].copy(), got[:, 1].copy()
    recv_off = np.zeros(W + 1, np.int64)
    np.cumsum(recv_bytes, out=recv_off[1:])
    return recv_bytes, recv_nrec, recv_off


def exchange_segments(send, send_off, recv_alloc, recv_bytes, group=None):
    """One variable-size all-to-all of the packed segments.
    send: uint8 tensor holding all outgoing segments; send_off: W+1 byte offsets;
    recv_alloc(nbytes) -> uint8 tensor to receive into."""
    send_sizes = np.diff(np.asarray(send_off, np.int64)).tolist()
    total = int(np.sum(recv_bytes))
    recv = recv_alloc(total)
    dist.all_to_all_single(recv[:total], send[:int(send_off[-1])], output_split_sizes=[int(x) for x in recv_bytes],
                           input_split_sizes=send_sizes, group=group)
    return recv


def shuffle(ctx, group=None):
    """mr_shuffle_export -> all-to-all -> mr_shuffle_import on a _native.Context (GPU ranks)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    base, off, nrec = ctx.shuffle_export()
    send = device_bytes(base, int(off[-1]), dev)
    recv_bytes, recv_nrec, recv_off = exchange_plan(np.diff(off.astype(np.int64)), nrec, group, dev)
    exchange_segments(send, off, lambda n: device_bytes(ctx.shuffle_recv_buffer(n), n, dev), recv_bytes, group)
    ctx.shuffle_import(recv_off.astype(np.uint64), recv_nrec.astype(np.uint64))
    return int(off[-1]), int(recv_off[-1])
