"""Row-strip planning for the multi-GPU driver: pure host logic, no device code.

Rank/device g of n owns the contiguous rows [g*H/n, (g+1)*H/n) and needs `halo` more rows on each side (clipped at
the true image border, where the kernels clamp).  The same split is used by csrc/mg.cu (vip_mg_strip_rows).
"""
from collections import namedtuple

from .capi import Strip

StripPlan = namedtuple("StripPlan", "y0 y1 s0 s1")  # owned rows [y0, y1); rows held incl. halo [s0, s1)


def halo_rows(kind, ksize, nitr=1, exchange_between_iterations=True):
    """Rows of halo one pass needs on each side.  bf/jbf/abf: radius.  btf: (k-1) + k/2 + k/2 + 1 per iteration; without
    an exchange between iterations the halos of all iterations add up (redundant recompute instead of communication)."""
    r = ksize // 2
    if kind in ("bf", "jbf", "abf"):
        return r
    if kind == "gradient":
        return 1
    if kind == "btf":
        per_it = (ksize - 1) + 2 * r + 1
        return per_it if exchange_between_iterations else per_it * max(1, nitr)
    raise ValueError(kind)


def plan_strips(height, n, halo):
    """One StripPlan per rank; every row is owned exactly once, ranks beyond `height` get empty strips."""
    if height <= 0 or n <= 0 or halo < 0:
        raise ValueError("plan_strips: bad arguments")
    plans = []
    for g in range(n):
        y0, y1 = height * g // n, height * (g + 1) // n
        s0, s1 = (max(0, y0 - halo), min(height, y1 + halo)) if y1 > y0 else (y0, y0)
        plans.append(StripPlan(y0, y1, s0, s1))
    return plans


def to_strip(plan):
    """struct vip_strip of a plan whose destination buffer holds exactly the owned rows."""
    return Strip(plan.s0, plan.s1 - plan.s0, plan.y0, plan.y0, plan.y1 - plan.y0)
