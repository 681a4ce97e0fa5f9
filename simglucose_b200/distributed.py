"""Multi-GPU plumbing: episodes are independent for their whole life, so a population shards by
contiguous episode ranges (one process per GPU) with no data-path collective.  The only exchange is
the final reduction of the population statistics vector (NCCL over NVLink on GPUs; gloo in the CPU
tests) -- the B200 counterpart of the reference's pathos map/gather (simulation/sim_engine.py:65-76).
"""
import torch
import torch.distributed as dist

from . import _capi

_MIN, _MAX = _capi.STAT_COLS.index('min_bg'), _capi.STAT_COLS.index('max_bg')


def shard_range(n_total, rank, world):
    """Contiguous [lo, hi) of rank's episodes; sizes differ by at most one."""
    base, rem = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_stats(agg, group=None):
    """In-place all-reduce of an aggregate statistics vector [NSTAT] (sg_reduce_stats layout): SUM for the
    additive entries, MIN / MAX for the BG extrema.  Three tiny latency-bound collectives."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return agg
    mn = agg[_MIN:_MIN + 1].clone()
    mx = agg[_MAX:_MAX + 1].clone()
    dist.all_reduce(agg, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    agg[_MIN] = mn[0]
    agg[_MAX] = mx[0]
    return agg


def summarize(agg):
    """Aggregate vector -> dict of population statistics (means over history rows)."""
    a = agg.detach().cpu().tolist() if isinstance(agg, torch.Tensor) else list(agg)
    s = dict(zip(_capi.STAT_COLS, a))
    rows = max(s['rows'], 1.0)
    return {
        'rows': s['rows'], 'mean_bg': s['sum_bg'] / rows, 'min_bg': s['min_bg'], 'max_bg': s['max_bg'],
        'mean_lbgi': s['sum_lbgi'] / rows, 'mean_hbgi': s['sum_hbgi'] / rows, 'mean_risk': s['sum_risk'] / rows,
        'sum_reward': s['sum_reward'],
        'pct_lt50': 100.0 * s['lt50'] / rows, 'pct_lt70': 100.0 * s['lt70'] / rows,
        'pct_in_range': 100.0 * s['inrange'] / rows, 'pct_gt180': 100.0 * s['gt180'] / rows,
        'pct_gt250': 100.0 * s['gt250'] / rows, 'episodes_done': s['first_done'],
        'dopri_steps': s['dopri_steps'], 'dopri_rejects': s['dopri_rejects'],
    }
