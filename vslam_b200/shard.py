"""Host-side sharding of an odometry sequence over ranks (one process per GPU).

The path partitions by frame pair: pair i aligns frame i+1 against frame i and depends on nothing else, so ranks
take contiguous pair ranges (each also needs the first frame of its range: one frame of overlap) and there is no
data-path collective.  Batches of independent pairs (BASELINE configs[2]) are dealt round-robin instead (deal_pairs).  The only exchange is the gather of the 7-double relative poses at the end
(torch.distributed all_gather: NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np


def pair_range(n_pairs, world, rank):
    """[lo, hi) of the pairs rank owns: contiguous, sizes differ by at most one, earlier ranks take the remainder."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad world/rank")
    base, rem = divmod(n_pairs, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def frame_range(n_pairs, world, rank):
    """[lo, hi) of the frames rank must hold: its pairs plus the leading keyframe (empty if it owns no pair)."""
    lo, hi = pair_range(n_pairs, world, rank)
    return (lo, hi + 1) if hi > lo else (lo, lo)


def deal_pairs(n_pairs, world, rank):
    """Indices of the pairs rank owns when independent pairs are dealt round-robin (pair k goes to rank k mod world).
    The alignments of a batch need different numbers of Gauss-Newton iterations; dealing them like cards evens the
    per-rank work out where contiguous blocks of a sorted or correlated batch would not."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad world/rank")
    return np.arange(rank, n_pairs, world)


def gather_dealt(local, n_pairs, dist=None, device="cpu"):
    """all_gather of per-rank (n_local, 7) pose arrays of round-robin dealt pairs (deal_pairs) into the (n_pairs, 7)
    array of the whole batch, in pair order, on every rank."""
    import torch
    local = np.ascontiguousarray(local, np.float64).reshape(-1, 7)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        if len(local) != n_pairs:
            raise ValueError("single rank must hold every pair")
        return local.copy()
    world, rank = dist.get_world_size(), dist.get_rank()
    if len(local) != len(deal_pairs(n_pairs, world, rank)):
        raise ValueError(f"rank {rank} holds {len(local)} poses, expected {len(deal_pairs(n_pairs, world, rank))}")
    cap = (n_pairs + world - 1) // world
    buf = torch.zeros((cap, 7), dtype=torch.float64, device=device)
    if len(local):
        buf[:len(local)] = torch.from_numpy(local).to(device)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.empty((n_pairs, 7))
    for r, p in enumerate(parts):
        idx = deal_pairs(n_pairs, world, r)
        out[idx] = p[:len(idx)].cpu().numpy()
    return out


def gather_poses(local, n_pairs, dist=None, device="cpu"):
    """all_gather of per-rank (n_local, 7) pose arrays into the (n_pairs, 7) array of the whole sequence, in pair
    order, on every rank.  Ragged shards are padded to the largest one for the collective."""
    import torch
    local = np.ascontiguousarray(local, np.float64).reshape(-1, 7)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        if len(local) != n_pairs:
            raise ValueError("single rank must hold every pair")
        return local.copy()
    world, rank = dist.get_world_size(), dist.get_rank()
    lo, hi = pair_range(n_pairs, world, rank)
    if len(local) != hi - lo:
        raise ValueError(f"rank {rank} holds {len(local)} poses, expected {hi - lo}")
    cap = max(pair_range(n_pairs, world, r)[1] - pair_range(n_pairs, world, r)[0] for r in range(world))
    buf = torch.zeros((cap, 7), dtype=torch.float64, device=device)
    if len(local):
        buf[:len(local)] = torch.from_numpy(local).to(device)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.empty((n_pairs, 7))
    for r, p in enumerate(parts):
        a, b = pair_range(n_pairs, world, r)
        out[a:b] = p[:b - a].cpu().numpy()
    return out


def quat_mul(a, b):
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz])


def quat_rot(q, v):
    qv = np.array([v[0], v[1], v[2], 0.0])
    qc = np.array([-q[0], -q[1], -q[2], q[3]])
    return quat_mul(quat_mul(q, qv), qc)[:3]


def pose_mul(a, b):
    """SE3 product of two poses in Sophus data() order (qx, qy, qz, qw, tx, ty, tz)."""
    q = quat_mul(a[:4], b[:4])
    return np.concatenate([q / np.linalg.norm(q), quat_rot(a[:4], b[4:]) + a[4:]])


def chain_trajectory(relative, first=None):
    """Absolute poses of frames 0..n from the relative motions pose_i = rel_i * pose_{i-1} (Odometry.cpp's
    frame-to-frame chaining: P_frame = pose * P_world)."""
    cur = np.array([0, 0, 0, 1, 0, 0, 0], np.float64) if first is None else np.asarray(first, np.float64)
    out = [cur]
    for r in np.asarray(relative, np.float64).reshape(-1, 7):
        cur = pose_mul(r, cur)
        out.append(cur)
    return np.stack(out)
