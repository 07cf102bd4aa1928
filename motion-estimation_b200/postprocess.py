"""Encoder ground truth, resampling and CSV output -- the host-side tail of the video script.

Reference: ``encoder_velocity`` estimate_vel_vid.py:417-487, ``dataset_correlation`` :589-622 and
the two CSV writers inside ``presentData`` :630-642 (the matplotlib half, :671-693, needs a display
and is left out).  Pure host code on lists of at most a few thousand rows; it is here so that the
drop-in produces every file the reference produces, not because it needs a GPU.

Two quirks of the reference are reproduced, because downstream users of its CSVs depend on them,
and flagged:
* ``velocity_estimates.csv``: the header says (timestamp, x, y) but the rows are (x, y, timestamp)
  (:410 vs :572);
* ``dataset_correlation`` returns "velocities" that are position / dt * 60, not differences (:612-613).
"""
from __future__ import annotations

import csv
import glob
import os
from typing import List, Sequence, Tuple

import numpy as np

VELOCITY_HEADER = ["Timestamp [s]", "mm/min X-axis optical", "mm/min Y-axis optical"]          # :410
POSITION_HEADER = ["timestamp [s]", "position optical", "position encoder"]                    # :628


def _first(directory, pattern):
    hits = sorted(glob.glob(os.path.join(directory, pattern)))
    if not hits:
        raise FileNotFoundError(f"no {pattern} in {directory}")
    return hits[0]


def encoder_velocity(imagePath: str, lastFrame: int, with_velocity: bool = False):
    """Encoder position / time series up to the optical recording's last timestamp (:417-487).

    The encoder log has one row per millisecond: column 0 = ms, column 2 = filament position in mm,
    column 8 = 1 once the extruder (and with it the camera trigger) is active (:441-456).  Time is
    re-based to the trigger row and accumulated gap by gap exactly as the reference does.
    Returns (enc_pos_list, enc_ts_list) [+ the (t, pos, mm/min) rows when ``with_velocity``].
    """
    stamp_lines = open(_first(imagePath, "*tstamps.txt")).read().split("\n")
    horizon_s = float(stamp_lines[lastFrame - 1]) / 1000.0            # linecache.getline(file, lastFrame), ms -> s (:429-433)
    rows = open(_first(imagePath, "*encoder.csv")).read().split("\n")

    def row(k):                                                        # 1-based like linecache; "" past the end
        return rows[k - 1] if 1 <= k <= len(rows) else ""

    k = 2
    while True:                                                        # first row whose trigger flag is set (:441-446)
        k += 1
        if int(row(k).split(",")[8].split("\n")[0][0]) != 0:
            break
    pos_list: List[float] = []
    ts_list: List[float] = []
    info: List[Tuple[float, float, float]] = []
    elapsed = 0.0
    prev_t = prev_p = None
    while elapsed <= horizon_s:                                        # :450
        line = row(k)
        if line == "":
            break
        cols = line.split(",")
        t_s, p = float(cols[0]) / 1000.0, float(cols[2])
        if prev_t is not None:
            gap = t_s - prev_t
            elapsed = elapsed + gap
            info.append((elapsed, p, (p - prev_p) / (gap / 60)))       # :470-471
            pos_list.append(p)
            ts_list.append(elapsed)
        prev_t, prev_p = t_s, p
        k += 1
    return (pos_list, ts_list, info) if with_velocity else (pos_list, ts_list)


def dataset_correlation(optical_pos: Sequence[float], optical_ts: Sequence[float], encoder_pos: Sequence[float],
                        encoder_ts: Sequence[float]):
    """Akima-resample both position series onto one uniform time grid (:589-622).

    Returns (interpolated_opt_pos, interpolated_enc_pos, tsList_new, vel_enc, vel_opt) like the
    reference; the two "velocity" lists keep its arithmetic (position / dt * 60, see module docstring).
    """
    from scipy import interpolate
    f_opt = interpolate.Akima1DInterpolator(optical_ts, optical_pos)
    f_enc = interpolate.Akima1DInterpolator(encoder_ts, encoder_pos)
    grid = np.linspace(0, optical_ts[-1], len(optical_ts))            # :599
    opt, enc = f_opt(grid), f_enc(grid)
    vel_opt, vel_enc = [], []
    for i in range(1, len(grid)):
        dt = grid[i] - grid[0]                                         # the reference never advances old_ts (:608-611)
        vel_opt.append((opt[i] / dt) * 60)
        vel_enc.append((enc[i] / dt) * 60)
    return opt, enc, grid, vel_enc, vel_opt


def write_velocity_csv(directory: str, out_information) -> str:
    """velocity_estimates.csv with the reference's header and row layout (:633-636)."""
    path = os.path.join(directory, "velocity_estimates.csv")
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(VELOCITY_HEADER)
        w.writerows(out_information)
    return path


def write_positions_csv(directory: str, grid, opt, enc) -> str:
    """correlated_positions.csv (:638-642)."""
    path = os.path.join(directory, "correlated_positions.csv")
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(POSITION_HEADER)
        for value in zip(grid, opt, enc):
            w.writerow(value)
    return path
