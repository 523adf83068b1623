"""The shuffle between ranks: one process per GPU, torch.distributed for the plumbing.

Replaces the reference's file-based all-to-all (every mapper writes <= R run files,
job.lua:208-221; every reducer lists and reads <= M of them, job.lua:253-260;
server.lua:279-329 discovers them).  Here every rank tallies its combined (key, count)
records per owner rank (owner = partition mod world_size, so each reduce partition is owned
by exactly one GPU), the tallies of all ranks are gathered once (the only host
synchronisation of the shuffle), every rank packs one segment per owner, the segments move
with ONE variable-size all-to-all (NCCL over NVLink on GPUs; gloo in the CPU tests of this
host logic), and the owner merges what it received with the radix sort + segmented reduce.

Segment wire format (per source -> destination): n records of 24 bytes {k0, k1, count}
followed by the bytes of the keys longer than 15 bytes, the whole rounded up to 16 bytes
(include/mrgpu.h, mr_shuffle_pack).
"""
import numpy as np
import torch
import torch.distributed as dist

from . import _native

REC_BYTES = 24


def shard_splits(n_splits, world_size, rank):
    """Split ids of one rank: round-robin, like workers claiming map jobs in turn (task.lua:258-343)."""
    return list(range(rank, n_splits, world_size))


def owner_of_partition(part, world_size):
    return part % world_size


class _DeviceBytes:
    """Wraps a raw device pointer so torch can view it (no copy)."""

    def __init__(self, ptr, n, typestr="|u1"):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


def device_bytes(ptr, n, device):
    if n == 0:
        return torch.empty(0, dtype=torch.uint8, device=device)
    return torch.as_tensor(_DeviceBytes(ptr, n), device=device)


def device_i64(ptr, n, device):
    return torch.as_tensor(_DeviceBytes(ptr, n, "<i8"), device=device)


def gather_counts(counts, group=None):
    """counts: int64 tensor [2W] of this rank (records per destination, then long-key bytes per
    destination), on the device the backend works with.  Returns the [W, 2W] matrix of all ranks
    as numpy -- the one host synchronisation of the shuffle."""
    W = dist.get_world_size(group)
    out = [torch.empty_like(counts) for _ in range(W)]
    dist.all_gather(out, counts.contiguous(), group=group)
    return torch.stack(out).cpu().numpy().astype(np.int64)


def segment_bytes(nrec, nbytes):
    return (nrec * REC_BYTES + nbytes + 15) // 16 * 16


def plan_from_counts(matrix, rank):
    """From the gathered counts: what this rank sends and receives.
    Returns (send_off [W+1], recv_off [W+1], recv_nrec [W]) as int64 numpy arrays."""
    W = matrix.shape[0]
    send = segment_bytes(matrix[rank, :W], matrix[rank, W:])
    recv = segment_bytes(matrix[:, rank], matrix[:, W + rank])
    send_off = np.zeros(W + 1, np.int64)
    recv_off = np.zeros(W + 1, np.int64)
    np.cumsum(send, out=send_off[1:])
    np.cumsum(recv, out=recv_off[1:])
    return send_off, recv_off, matrix[:, rank].copy()


def exchange_segments(send, send_off, recv, recv_off, group=None):
    """One variable-size all-to-all of the packed segments (uint8 tensors)."""
    dist.all_to_all_single(recv[:int(recv_off[-1])], send[:int(send_off[-1])],
                           output_split_sizes=np.diff(recv_off).tolist(),
                           input_split_sizes=np.diff(send_off).tolist(), group=group)
    return recv


def shuffle(ctx, group=None, slot_bytes=None, stats=None):
    """The shuffle of one job on a _native.Context (GPU ranks).  stats: a dict that receives the
    size matrix of an exact-size shuffle ("matrix": [src][records per dst..., long-key bytes per dst...]).

    slot_bytes=None: exact sizes -- count -> gather (one host sync) -> pack -> all-to-all -> import.
        Returns the largest segment any rank sends to any rank, which a caller running the same
        job shape again can turn into a slot size.
    slot_bytes=N:    fixed slots, sizes in device-side headers, equal-split all-to-all, no host
        synchronisation at all; a segment larger than N surfaces as MR_ERR_OVERFLOW at job_sync."""
    dev = torch.device("cuda", torch.cuda.current_device())
    W, rank = ctx.world_size, ctx.rank
    # Stream contract (include/mrgpu.h, shuffle seam): count / pack / import are asynchronous on the
    # context's stream, and the collectives below are ordered against torch's current stream only --
    # so the two must be the same stream, or the all-gather could read the tallies before the count
    # kernel wrote them and the all-to-all could move a send buffer that is not packed yet.
    cur = torch.cuda.current_stream(dev).cuda_stream
    if getattr(ctx, "stream", cur) != cur:
        raise _native.NativeError(_native.MR_ERR_INVALID,
                                  "parallel.shuffle: the context must run on torch's current stream "
                                  "(create it with stream=torch.cuda.current_stream().cuda_stream on a "
                                  "non-default stream); a context with a private stream is not ordered "
                                  "against the collectives")
    if slot_bytes:
        total = W * int(slot_bytes)
        send = device_bytes(ctx.shuffle_pack_fixed(int(slot_bytes)), total, dev)
        recv = device_bytes(ctx.shuffle_recv_buffer(total), total, dev)
        dist.all_to_all_single(recv, send, group=group)
        ctx.shuffle_import_fixed(int(slot_bytes))
        return int(slot_bytes)
    counts = device_i64(ctx.shuffle_count(), 2 * W, dev)
    matrix = gather_counts(counts, group)
    if stats is not None:
        stats["matrix"] = matrix
    send_off, recv_off, recv_nrec = plan_from_counts(matrix, rank)
    base, off, nrec = ctx.shuffle_pack(matrix[rank].astype(np.uint64))
    assert np.array_equal(off.astype(np.int64), send_off)
    send = device_bytes(base, int(send_off[-1]), dev)
    recv = device_bytes(ctx.shuffle_recv_buffer(int(recv_off[-1])), int(recv_off[-1]), dev)
    exchange_segments(send, send_off, recv, recv_off, group)
    ctx.shuffle_import(recv_off.astype(np.uint64), recv_nrec.astype(np.uint64))
    return int(segment_bytes(matrix[:, :W], matrix[:, W:]).max())


def split_rows(part_counts, world):
    """Where the world-1 equal-share cut points of a finished job fall.  part_counts[r][p] = rows of
    partition p held by rank r (within a partition the ranks' slices follow each other in rank
    order, under either ownership rule).  Returns [(rank, row on that rank)] for cuts 1..world-1,
    or [] for an empty job."""
    c = np.asarray(part_counts, dtype=np.int64)
    total = int(c.sum())
    if total == 0:
        return []
    begins = np.concatenate([np.zeros((c.shape[0], 1), np.int64), np.cumsum(c, axis=1)], axis=1)  # per-rank part_begin
    blocks = [(p, r) for p in range(c.shape[1]) for r in range(c.shape[0]) if c[r, p]]             # global output order
    ends = np.cumsum([c[r, p] for p, r in blocks])
    cuts = []
    for j in range(1, world):
        pos = j * total // world
        b = int(np.searchsorted(ends, pos, side="right"))
        p, r = blocks[b]
        cuts.append((r, int(begins[r, p] + pos - (ends[b] - c[r, p]))))
    return cuts


def balance_owners(ctx, group=None):
    """Collective, after a finished job on every rank: installs world-1 splitters at equal shares
    of that job's rows, so that later jobs of the same shape give every rank the same number of
    keys to merge (mr_shuffle_set_owners).  Returns the splitters."""
    info = ctx.job_sync()
    rows = ctx.rows_part_begin()
    counts = [None] * ctx.world_size
    dist.all_gather_object(counts, np.diff(rows.astype(np.int64)).tolist(), group=group)
    cuts = split_rows(counts, ctx.world_size)
    mine = {j: ctx.row_sort_word(row) for j, (r, row) in enumerate(cuts) if r == ctx.rank}
    parts = [None] * ctx.world_size
    dist.all_gather_object(parts, mine, group=group)
    merged = {}
    for d in parts:
        merged.update(d)
    splitters = [merged[j] for j in range(len(cuts))]
    ctx.set_owners(splitters)
    return splitters


def p2p_slot_for(max_segment_bytes, headroom=1.25):
    """Slot size for the peer-to-peer shuffle from the largest exact segment seen: its records are
    32 bytes where the packed segments have 24, its header 32 where they have 16."""
    return (int(max_segment_bytes * headroom * 4 / 3) + 32 + 4095) // 4096 * 4096


def p2p_connect(ctx, slot_bytes, group=None):
    """Collective: every rank sets up its receive window and maps everybody else's (the 128-byte
    blobs travel through the process group, which stands for the host's control plane).  Where any
    rank cannot map a peer's memory, EVERY rank raises NativeError (so that all of them fall back to
    another transport together)."""
    mine = ctx.p2p_setup(int(slot_bytes))
    blobs = [None] * ctx.world_size
    dist.all_gather_object(blobs, mine, group=group)
    err = None
    try:
        ctx.p2p_attach(blobs)
    except _native.NativeError as e:
        err = e
    ok = [None] * ctx.world_size
    dist.all_gather_object(ok, err is None, group=group)   # also: nobody signals into a window that is not there yet
    if not all(ok):
        raise err or _native.NativeError(_native.MR_ERR_UNSUPPORTED, "a peer could not map the receive windows")
    return int(slot_bytes)


def slot_for(max_segment_bytes, headroom=1.25):
    """A slot size for the fixed-slot shuffle from the largest exact segment seen (header included)."""
    return (int(max_segment_bytes * headroom) + 16 + 4095) // 4096 * 4096
