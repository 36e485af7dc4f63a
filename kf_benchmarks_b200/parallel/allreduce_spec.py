"""The ``--all_reduce_spec`` grammar and gradient-bucket planning helpers.

Grammar (T/benchmark_cnn.py:536-553, parser T/allreduce.py:58-104)::

    spec  ::= range | range ':' size ':' spec
    range ::= alg [ '#' shards ]
    alg   ::= nccl | nccl/xring | nccl/rechd | nccl/pscpu | xring | pscpu | psgpu |
              pscpu/pscpu | collective
    size  ::= int [ K | M | G | T ]         (upper size limit of the tensors the range covers)

On an NVSwitch box every algorithm name selects the same fused P2P kernel family (one-shot
for small buckets, two-shot above); the size limits still decide the bucket boundaries.
"""
from __future__ import annotations

import re
from collections import namedtuple
from typing import List, Sequence

AllReduceSpecTuple = namedtuple("AllReduceSpecTuple", "alg shards limit")

_ALGS = ("nccl", "nccl/xring", "nccl/rechd", "nccl/pscpu", "xring", "pscpu", "psgpu",
         "pscpu/pscpu", "collective")


def parse_general_int(s: str) -> int:
    """'64' -> 64, '4k' -> 4096, '2M' -> 2097152 ... (T/allreduce.py:35-55)."""
    m = re.match(r"^(\d+)([KkMGT]?)$", s)
    if not m:
        raise ValueError("invalid integer representation %s" % s)
    mult = {"": 1, "K": 1 << 10, "k": 1 << 10, "M": 1 << 20, "G": 1 << 30, "T": 1 << 40}
    return int(m.group(1)) * mult[m.group(2)]


def parse_all_reduce_spec(all_reduce_spec: str) -> List[AllReduceSpecTuple]:
    range_parts = all_reduce_spec.split(":") + ["-1"]
    if len(range_parts) % 2:
        raise ValueError("all_reduce_spec not well formed: %s" % all_reduce_spec)
    limit = 0
    spec = []
    for i in range(0, len(range_parts), 2):
        alg_part, limit_part = range_parts[i], range_parts[i + 1]
        if "#" in alg_part:
            alg, shards = alg_part.split("#")
            shards = int(shards)
        else:
            alg, shards = alg_part, 1
        if alg not in _ALGS:
            raise ValueError("all_reduce_spec (%s) contains invalid alg %s" %
                             (all_reduce_spec, alg))
        if limit_part == "-1":
            new_limit = -1
        else:
            new_limit = parse_general_int(limit_part)
            if new_limit < limit:
                raise ValueError("all_reduce_spec (%s) contains non-ascending limit" %
                                 all_reduce_spec)
            limit = new_limit
        spec.append(AllReduceSpecTuple(alg=alg, shards=shards, limit=new_limit))
    return spec


def extract_ranges(index_list: Sequence[int], range_size_limit: int = 32):
    """Splits sorted indices into maximal consecutive runs (returned as (first, last)) of at
    most ``range_size_limit`` and leftover singles (T/allreduce.py:420-454)."""
    if not index_list:
        return [], []
    first = last = index_list[0]
    ranges, singles = [], []
    for i in index_list[1:]:
        if i == last + 1 and (last - first) <= range_size_limit:
            last = i
        else:
            if last > first:
                ranges.append([first, last])
            else:
                singles.append(first)
            first = last = i
    if last > first:
        ranges.append([first, last])
    else:
        singles.append(first)
    return ranges, singles


def small_tensor_groups(sizes_bytes: Sequence[int], max_bytes: int, max_group: int):
    """Which consecutive small tensors would be folded together (--agg_small_grads_*):
    returns (groups of indices, indices left alone). With the flat arena this only decides
    bucket boundaries - nothing is copied."""
    small = [i for i, s in enumerate(sizes_bytes) if s <= max_bytes]
    ranges, singles = extract_ranges(small, range_size_limit=max_group)
    groups = [list(range(a, b + 1)) for a, b in ranges]
    return groups, singles
