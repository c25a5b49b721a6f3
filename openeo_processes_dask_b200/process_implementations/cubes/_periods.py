"""Calendar bins for ``aggregate_temporal_period`` (host logic, pandas only).

* hour/day/week/month/season/year: the reference resamples with the pandas frequencies
  ``h, D, W, ME, QS-DEC, YS`` (``cubes/aggregate.py:231-247``); xarray's resample is a pandas Grouper
  with pandas' default ``closed``/``label`` per frequency, empty bins included -> restated with
  ``Series.resample``.
* dekad/decade/decade-ad/tropical-season: the reference synthesises interval bounds and labels
  (``get_intervals``, ``cubes/aggregate.py:104-203``) and calls ``aggregate_temporal``.  ``period_bounds``
  below produces the SAME bounds and labels as that function does when it is run from the reference source
  (tests/golden/reference_cubes.json: 569 extents, checked case by case in tests/test_cubes_golden.py),
  quirks included, because a drop-in has to return the labels the reference returns:
    - decades: Jan-1 of floor(y0/10)*10 .. ceil(y1/10)*10 in steps of 10 years (+1 for decade-ad); an extent
      inside one year divisible by 10 gives ONE bound, i.e. no interval at all;
    - tropical seasons: the bounds are the LAST day of November / May (``period_range(freq="6M").strftime``
      prints period ends), labels ``YYYY-ndjfma`` / ``YYYY-mjjaso`` from the bound's own year;
    - dekads: the first bound is day ``floor(day/10)*10 + 1`` of the first month (day 10 -> 11, day 20 -> 21,
      day 30 -> 31, which does not exist in a 30-day month: the reference raises, so does this), whole months in
      between contribute 1 / 11 / 21, the last month is appended again from its 1st -- for an extent inside one
      month the bound list is therefore not monotonic and ``aggregate_temporal`` rejects it as overlapping, as
      in the reference; labels are ``YYYY-<int(day_of_year / 10)>``.
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
    return np.array([[a, b] for a, b in zip(text[:-1], text[1:])], dtype=str).reshape(-1, 2)


def _month_end(year, month):
    return pd.Timestamp(year=year, month=month, day=1) + pd.offsets.MonthEnd(0)


def _decade_bounds(y0, y1, shift):
    lo = int(np.floor(y0 / 10) * 10) + shift
    hi = int(np.ceil(y1 / 10) * 10) + shift
    bounds = [pd.Timestamp(year=y, month=1, day=1) for y in range(lo, hi + 1, 10)]
    return bounds, [f"{b.year}" for b in bounds[:-1]]


def _tropical_bounds(start, end):
    # half-year periods labelled by their last month: (y, 5) = Dec..May, (y, 11) = Jun..Nov; printed as month ends
    if 5 <= start.month < 10:
        y, m = start.year, 5
    elif start.month < 5:
        y, m = start.year - 1, 11
    else:
        y, m = start.year, 11
    if 5 <= end.month < 10:
        stop = (end.year, 11)
    elif end.month < 5:
        stop = (end.year, 5)
    else:
        stop = (end.year + 1, 5)
    bounds = []
    while (y, m) <= stop:
        bounds.append(_month_end(y, m))
        y, m = (y, 11) if m == 5 else (y + 1, 5)
    labels = [f"{b.year}-" + ("ndjfma" if b.month == 11 else "mjjaso") for b in bounds[:-1]]
    return bounds, labels


def _dekad_bounds(start, end):
    first_day = int(np.floor(start.day / 10) * 10 + 1)
    # an impossible date (day 31 of a 30-day month) raises here like the reference's pd.date_range does
    first = pd.Timestamp(f"{start.year}-{start.month}-{first_day}")
    bounds = [first + pd.Timedelta(days=10 * k) for k in range(3) if (first + pd.Timedelta(days=10 * k)).day <= 22
              and (first + pd.Timedelta(days=10 * k)).month == first.month]
    # whole months strictly after the first one; the month `end` lies in is appended separately below
    anchor = pd.Timestamp(year=start.year, month=start.month, day=22)
    month_starts = list(pd.date_range(start=anchor, end=end, freq="1MS"))[:-1]
    for ms in month_starts:
        bounds += [ms, ms + pd.Timedelta(days=10), ms + pd.Timedelta(days=20)]
    last_first = pd.Timestamp(year=end.year, month=end.month, day=1)
    n = int(np.ceil((end.day - 1) / 10))
    if end.day > 21:
        bounds += [last_first + pd.Timedelta(days=10 * k) for k in range(n)]
        bounds.append(last_first + pd.offsets.MonthBegin(1))
    else:
        bounds += [last_first + pd.Timedelta(days=10 * k) for k in range(n + 1)]
    labels = []
    for b in bounds[:-1]:
        k = int(b.day_of_year / 10)
        labels.append(f"{b.year}-{k}" if k > 9 else f"{b.year}-0{k}")
    return bounds, labels


def period_bounds(t_values, period):
    t = np.asarray(t_values).astype("datetime64[ns]")
    start, end = pd.Timestamp(t[0]), pd.Timestamp(t[-1])
    if period == "decade":
        return _decade_bounds(start.year, end.year, 0)
    if period == "decade-ad":
        return _decade_bounds(start.year, end.year, 1)
    if period == "tropical-season":
        return _tropical_bounds(start, end)
    if period == "dekad":
        return _dekad_bounds(start, end)
    raise NotImplementedError(f"The provided period '{period})' is not implemented. ")


def period_intervals(t_values, period):
    """(intervals [[start, end], ...] as strings, labels) like the reference's ``get_intervals``."""
    bounds, labels = period_bounds(t_values, period)
    return _pairs(bounds), list(labels)
