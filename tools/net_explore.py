#!/usr/bin/env python3
"""Op counts of pruned comparator networks for the median wires (exploration tool, not part of the build)."""
import sys, random
sys.path.insert(0, __file__.rsplit('/', 1)[0])
from gen_sort_networks import batcher, apply


def prune_ops(cs, n, outs):
    """ops after dead-output elimination: a comparator with both outputs live = 2 ops, one live = 1."""
    live = set(outs)
    ops = 0; kept = []
    for a, b in reversed(cs):
        la, lb = a in live, b in live
        if not (la or lb):
            continue
        ops += la + lb
        kept.append((a, b, la, lb))
        live.add(a); live.add(b)
    kept.reverse()
    return ops, kept


def oe_merge(lo_idx, hi_idx):
    """Batcher odd-even merge of two sorted wire lists (arbitrary lengths) -> comparators; result sorted on
    the concatenated wire list order lo_idx + hi_idx positions."""
    cs = []
    def merge(a, b):  # a, b lists of wires, each sorted ascending; returns list of wires in sorted order
        if not a: return b
        if not b: return a
        if len(a) == 1 and len(b) == 1:
            cs.append((a[0], b[0])); return [a[0], b[0]]
        ev = merge(a[0::2], b[0::2])
        od = merge(a[1::2], b[1::2])
        # interleave: result[0]=ev[0], then compare od[i] with ev[i+1]
        res = [ev[0]]
        i = 0
        # standard: for general lengths, pair od[i] with ev[i+1]
        while i < len(od) or i + 1 < len(ev):
            if i < len(od) and i + 1 < len(ev):
                cs.append((od[i], ev[i + 1]) if True else None)
                res += [od[i], ev[i + 1]]
            elif i < len(od):
                res.append(od[i])
            else:
                res.append(ev[i + 1])
            i += 1
        return res
    out = merge(lo_idx, hi_idx)
    return cs, out


if __name__ == '__main__':
    n = 100
    cs = batcher(n)
    ops, _ = prune_ops(cs, n, [49, 50])
    print('batcher100 median ops', ops, 'full', 2 * len(cs))
    for m in (25, 50):
        c = batcher(m); print('batcher', m, len(c))
