"""Calendar bins for ``aggregate_temporal_period`` (host logic, pandas only).

* hour/day/week/month/season/year: the reference resamples with the pandas frequencies
  ``h, D, W, ME, QS-DEC, YS`` (``cubes/aggregate.py:231-247``); xarray's resample is a pandas Grouper
  with pandas' default ``closed``/``label`` per frequency, empty bins included -> restated with
  ``Series.resample``.
* dekad/decade/decade-ad/tropical-season: the reference synthesises intervals and labels
  (``cubes/aggregate.py:104-203``) and calls ``aggregate_temporal``; the intervals are rebuilt here
  from the period definitions of the openEO spec (``aggregate_temporal_period.json``), with the
  reference's labels: ``YYYY`` for decades, ``YYYY-ndjfma`` / ``YYYY-mjjaso`` for tropical seasons,
  ``YYYY-DD`` (dekad index 00..35) for dekads.
"""

from __future__ import annotations

import numpy as np
import pandas as pd

_FMT = "%Y-%m-%dT%H:%M:%S"


def calendar_groups(t_values, freq):
    idx = pd.DatetimeIndex(np.asarray(t_values).astype("datetime64[ns]"))
    series = pd.Series(np.arange(len(idx)), index=idx)
    groups, labels = [], []
    for label, members in series.resample(freq):
        groups.append([int(i) for i in members.values])
        labels.append(label.to_datetime64())
    return groups, np.asarray(labels, dtype="datetime64[ns]")


def _pairs(bounds):
    text = [pd.Timestamp(b).strftime(_FMT) for b in bounds]
    return np.array([[a, b] for a, b in zip(text[:-1], text[1:])], dtype=str)


def _dekad_start(ts: pd.Timestamp) -> pd.Timestamp:
    first = 1 if ts.day <= 10 else (11 if ts.day <= 20 else 21)
    return pd.Timestamp(year=ts.year, month=ts.month, day=first)


def _next_dekad(ts: pd.Timestamp) -> pd.Timestamp:
    if ts.day < 21:
        return ts + pd.Timedelta(days=10)
    return (ts + pd.offsets.MonthBegin(1)).normalize()


def period_intervals(t_values, period):
    t = np.asarray(t_values).astype("datetime64[ns]")
    start, end = pd.Timestamp(t[0]), pd.Timestamp(t[-1])
    if period in ("decade", "decade-ad"):
        shift = 1 if period == "decade-ad" else 0
        first = int(np.floor(start.year / 10) * 10) + shift
        last = int(np.ceil(end.year / 10) * 10) + shift
        bounds = [pd.Timestamp(year=y, month=1, day=1) for y in range(first, last + 1, 10)]
        labels = [f"{b.year:04d}" for b in bounds[:-1]]
    elif period == "tropical-season":
        if 5 <= start.month < 10:
            first = pd.Timestamp(year=start.year, month=5, day=1)
        elif start.month < 5:
            first = pd.Timestamp(year=start.year - 1, month=11, day=1)
        else:
            first = pd.Timestamp(year=start.year, month=11, day=1)
        if 5 <= end.month < 10:
            last = pd.Timestamp(year=end.year, month=11, day=1)
        elif end.month < 5:
            last = pd.Timestamp(year=end.year, month=5, day=1)
        else:
            last = pd.Timestamp(year=end.year + 1, month=5, day=1)
        bounds = [first]
        while bounds[-1] < last:
            bounds.append(bounds[-1] + pd.DateOffset(months=6))
        labels = [f"{b.year:04d}-" + ("ndjfma" if b.month == 11 else "mjjaso") for b in bounds[:-1]]
    elif period == "dekad":
        bounds = [_dekad_start(start)]
        stop = _next_dekad(_dekad_start(end))
        while bounds[-1] < stop:
            bounds.append(_next_dekad(bounds[-1]))
        labels = [f"{b.year:04d}-{(b.month - 1) * 3 + (b.day - 1) // 10:02d}" for b in bounds[:-1]]
    else:
        raise NotImplementedError(f"The provided period '{period})' is not implemented. ")
    return _pairs(bounds), list(labels)
