"""Pair loop and velocity reduction -- host mirror of the reference's ``end_process``.

Reference: estimate_vel_vid.py:493-583 (anchor-frame loop of the video script) and
estimate_vel.py:333-378 (strictly consecutive loop of the raw-image script).  The arithmetic on
top of a pair displacement is a handful of float64 operations per accepted pair, so it stays on
the host and follows the reference statement by statement; the displacement itself comes from a
``PairSource`` (the GPU engine in the product, anything with the same two methods in tests).

The video script's loop is sequentially dependent: frame k is compared with the last ACCEPTED
frame, not with k-1 (:531-534, :575).  ``AnchoredPairs`` keeps that rule exact while still
batching: it asks the source for (anchor, k), (anchor, k+1), ... in one launch, walks the results
until the first accepted k, and throws the speculated remainder away.  d(anchor, k) is the same
computation whether it was speculated or not, so the output is identical to the serial loop.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

# estimate_vel_vid.py:396-399 / estimate_vel.py:308-312
PIXELS_PER_MM = 611
MAX_FILAMENT_SPEED_MM_MIN = 140


@dataclass
class PairLoopResult:
    """What end_process returns (estimate_vel_vid.py:583), plus the bookkeeping it only prints."""
    out_information: List[Tuple[float, float, float]] = field(default_factory=list)   # (vxm, vym, t_s)
    velocity_list_x: List[float] = field(default_factory=list)
    velocity_list_y: List[float] = field(default_factory=list)
    tsList: List[float] = field(default_factory=list)
    motion_list_opt: List[float] = field(default_factory=list)
    failed_estimates: int = 0                 # :536-537
    accepted_pairs: List[Tuple[int, int]] = field(default_factory=list)   # (anchor index, frame index)
    pair_displacements: List[Tuple[float, float, int]] = field(default_factory=list)
    searches: int = 0                         # pair searches actually run (speculation included)
    launches: int = 0
    # the reference's second opinion (:524-526): the frame mean of the dense field of EVERY compared pair
    # (anchor, k), accepted or not, summed in full-resolution pixels; :582 prints it divided by pixels_per_mm
    total_motion_optical_flow: float = 0.0
    opt_flow_list: List[float] = field(default_factory=list)
    final_position_dense: float = 0.0

    def as_tuple(self):
        return (self.out_information, self.velocity_list_x, self.velocity_list_y, self.tsList, self.motion_list_opt)


class AnchoredPairs:
    """Exact evaluation of d(anchor, k) with look-ahead batching.

    ``search_batch(anchor, ks)`` must return one (pdx, pdy, flag) per k in ``ks`` -- displacement of
    frame k relative to frame ``anchor`` in the units the gate below expects.
    """

    def __init__(self, search_batch: Callable[[int, Sequence[int]], Sequence[Tuple[float, float, int]]],
                 n_frames: int, min_lookahead: int = 4, max_lookahead: int = 256):
        self._search = search_batch
        self._n = n_frames
        self._lo, self._hi = min_lookahead, max_lookahead
        self._look = min_lookahead
        self._anchor = -1
        self._cache = {}
        self.searches = 0
        self.launches = 0

    def get(self, anchor: int, k: int):
        if anchor != self._anchor:
            # a new anchor was accepted: size the next speculation from how far the last one went
            self._cache.clear()
            self._anchor = anchor
            self._look = self._lo
        if k not in self._cache:
            ks = list(range(k, min(k + self._look, self._n)))
            res = self._search(anchor, ks)
            self.searches += len(ks)
            self.launches += 1
            for kk, r in zip(ks, res):
                self._cache[kk] = r
            self._look = min(self._look * 2, self._hi)   # still the same anchor next time: look further
        return self._cache[k]


def anchored_pair_loop(n_frames: int, pair: Callable[[int, int], Tuple[float, float, int]],
                       timestamp_ms: Callable[[int], float], first_frame: int = 0,
                       pixels_per_mm: float = PIXELS_PER_MM, gate_px: float = 2.0) -> PairLoopResult:
    """The video script's loop (estimate_vel_vid.py:493-583) over frames 0..n_frames-1.

    ``pair(a, k)`` -> (pdx, pdy, measurement_flag) or (pdx, pdy, measurement_flag, dense_mean_x) -- the fourth
    value is what ``calculate_dense_flow`` returns first at :524; ``timestamp_ms(line_no)`` returns the value on
    1-based line ``line_no`` of the timestamp file, as ``linecache.getline`` does at :540-541.
    """
    out = PairLoopResult()
    timestamp_k = first_frame + 2            # :498 "skip lines with metadata and first timestamp"
    old_ts = 0.0
    position = 0.0
    anchor = 0
    timestamp_k += 1                         # :511, first frame only becomes the anchor
    for k in range(1, n_frames):
        r = pair(anchor, k)                                               # :522
        pdx, pdy, flag = r[0], r[1], r[2]
        if len(r) > 3:                                                     # :524-526, before the gate
            out.total_motion_optical_flow += float(r[3])
            out.opt_flow_list.append(out.total_motion_optical_flow)
        if pdx < gate_px and flag == 1:                                    # :531-534 anchor kept
            timestamp_k += 1
            continue
        if flag == 0:                                                      # :536-537
            out.failed_estimates += 1
        ts = float(timestamp_ms(timestamp_k)) / 1000.0                     # :540-544
        gap_s = ts - old_ts                                                # :545
        gap_m = gap_s / 60.0                                               # :546
        out.tsList.append(ts)                                              # :548
        old_ts = ts                                                        # :551
        mm_dx, mm_dy = pdx / pixels_per_mm, pdy / pixels_per_mm            # :556
        vxm, vym = mm_dx / gap_m, mm_dy / gap_m                            # :560
        position += mm_dx                                                  # :562
        out.velocity_list_x.append(vxm)                                    # :568-570
        out.velocity_list_y.append(vym)
        out.motion_list_opt.append(position)
        out.out_information.append((vxm, vym, ts))                         # :572-573
        out.accepted_pairs.append((anchor, k))
        out.pair_displacements.append((float(pdx), float(pdy), int(flag)))
        anchor = k                                                         # :575
        timestamp_k += 1                                                   # :577
    out.final_position_dense = out.total_motion_optical_flow / pixels_per_mm   # :582
    return out


def consecutive_pair_loop(displacements: Sequence[Tuple[float, float]], timestamps_us: Sequence[float],
                          pixels_per_mm: float = PIXELS_PER_MM) -> PairLoopResult:
    """The raw-image script's loop (estimate_vel.py:344-378): pair k is (frame k-1, frame k).

    ``timestamps_us[k]`` is column 0 of line ``numberList[k]`` of tstamps.csv (:352-353).  The
    reference divides that microsecond value by 10E+6 (= 1e7, :354) and uses the absolute time, not
    the gap, as the divisor (:365-366); both are reproduced as they are.
    """
    out = PairLoopResult()
    for k, (pdx, pdy) in enumerate(displacements, start=1):
        t_s = int(timestamps_us[k]) / (10E+6)                              # :354
        t_min = t_s / 60                                                   # :355
        out.tsList.append(t_s)
        mm_dx, mm_dy = pdx / pixels_per_mm, pdy / pixels_per_mm            # :362
        vxm, vym = mm_dx / t_min, mm_dy / t_min                            # :366
        out.velocity_list_x.append(vxm)
        out.velocity_list_y.append(vym)
        out.out_information.append((vxm, vym, t_s))                        # :374
        out.accepted_pairs.append((k - 1, k))
        out.pair_displacements.append((float(pdx), float(pdy), 1))
    return out


def xmax_px(gap_minutes: float, pixels_per_mm: float = PIXELS_PER_MM,
            max_speed_mm_min: float = MAX_FILAMENT_SPEED_MM_MIN) -> float:
    """The printed-only physical bound of estimate_vel_vid.py:399,564: px the filament can move in a gap."""
    return pixels_per_mm * max_speed_mm_min * gap_minutes


def read_timecodes(path: str) -> List[str]:
    """Raw lines of a ``# timecode format v2`` file (1-based access like linecache: line 1 is the header)."""
    with open(path, "r") as f:
        return f.read().split("\n")


def timecode_lookup(lines: Sequence[str]) -> Callable[[int], float]:
    """``linecache.getline(file, n)`` followed by float() as at estimate_vel_vid.py:540-541."""
    def get(line_no: int) -> float:
        if line_no < 1 or line_no > len(lines) or lines[line_no - 1].strip() == "":
            raise ValueError(f"timestamp file has no value on line {line_no} (the reference raises here too)")
        return float(lines[line_no - 1])
    return get
