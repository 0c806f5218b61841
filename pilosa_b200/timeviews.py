"""Time-quantum views of a time field (host-side naming logic only; the bits live in ordinary fragments).

A `Row(f=r, from=..., to=...)` on a time field is the union of row r over the views that tile the time range
(executor.go:1501-1531): `views_by_time_range` restates viewsByTimeRange (time.go:104-181) with Go's
`time.AddDate` normalisation (Jan 31 + 1 month = Mar 2/3), pinned by the table of TestViewsByTimeRange
(time_internal_test.go:87-160). The union itself is an ordinary n-ary Union of Row leaves, one per view family,
so the engine needs nothing new for it.
"""
from __future__ import annotations

from datetime import datetime, timedelta
from typing import List

VIEW_STANDARD = "standard"


def _add_date(t: datetime, years: int, months: int, days: int) -> datetime:
    """Go's Time.AddDate: add to the fields, then normalise (day overflow runs into the next month)."""
    m = t.month - 1 + months
    y = t.year + years + m // 12
    m = m % 12 + 1
    return datetime(y, m, 1, t.hour, t.minute, t.second) + timedelta(days=t.day - 1 + days)


def _add_month(t: datetime) -> datetime:  # time.go:183-189
    if t.day > 28:
        t = datetime(t.year, t.month, 1, t.hour)
    return _add_date(t, 0, 1, 0)


def _next_year_gte(t: datetime, end: datetime) -> bool:  # time.go:191-197
    nxt = _add_date(t, 1, 0, 0)
    return nxt.year == end.year or end > nxt


def _next_month_gte(t: datetime, end: datetime) -> bool:  # time.go:199-207
    nxt = _add_date(t, 0, 1, 0)
    return (nxt.year, nxt.month) == (end.year, end.month) or end > nxt


def _next_day_gte(t: datetime, end: datetime) -> bool:  # time.go:209-217
    nxt = _add_date(t, 0, 0, 1)
    return (nxt.year, nxt.month, nxt.day) == (end.year, end.month, end.day) or end > nxt


def view_by_time_unit(name: str, t: datetime, unit: str) -> str:  # time.go:75-88
    fmt = {"Y": "%Y", "M": "%Y%m", "D": "%Y%m%d", "H": "%Y%m%d%H"}.get(unit)
    return f"{name}_{t.strftime(fmt)}" if fmt else ""


def views_by_time(name: str, t: datetime, quantum: str) -> List[str]:  # time.go:91-101 (the views a timestamped bit is written to)
    return [v for v in (view_by_time_unit(name, t, u) for u in quantum) if v]


def views_by_time_range(name: str, start: datetime, end: datetime, quantum: str) -> List[str]:  # time.go:104-181
    t = start
    has_y, has_m, has_d, has_h = ("Y" in quantum), ("M" in quantum), ("D" in quantum), ("H" in quantum)
    out: List[str] = []
    if has_h or has_d or has_m:  # walk up from the smallest unit to the largest
        while t < end:
            if has_h:
                if not _next_day_gte(t, end):
                    break
                if t.hour != 0:
                    out.append(view_by_time_unit(name, t, "H"))
                    t += timedelta(hours=1)
                    continue
            if has_d:
                if not _next_month_gte(t, end):
                    break
                if t.day != 1:
                    out.append(view_by_time_unit(name, t, "D"))
                    t = _add_date(t, 0, 0, 1)
                    continue
            if has_m:
                if not _next_year_gte(t, end):
                    break
                if t.month != 1:
                    out.append(view_by_time_unit(name, t, "M"))
                    t = _add_month(t)
                    continue
            break
    while t < end:  # and back down from the largest to the smallest
        if has_y and _next_year_gte(t, end):
            out.append(view_by_time_unit(name, t, "Y"))
            t = _add_date(t, 1, 0, 0)
        elif has_m and _next_month_gte(t, end):
            out.append(view_by_time_unit(name, t, "M"))
            t = _add_month(t)
        elif has_d and _next_day_gte(t, end):
            out.append(view_by_time_unit(name, t, "D"))
            t = _add_date(t, 0, 0, 1)
        elif has_h:
            out.append(view_by_time_unit(name, t, "H"))
            t += timedelta(hours=1)
        else:
            break
    return out


def view_time_part(v: str) -> str:  # time.go:334-339: "standard_201901" -> "201901"
    return v.split("_")[-1]


def min_max_views(views, quantum: str):  # time.go:240-274
    """The first and the last view (in string order) whose time part has the length of the quantum's LARGEST unit;
    ("", "") when there is none."""
    vs = sorted(views)
    chars = 4 if "Y" in quantum else 6 if "M" in quantum else 8 if "D" in quantum else 10 if "H" in quantum else 0
    hit = [v for v in vs if len(view_time_part(v)) == chars]
    return (hit[0], hit[-1]) if hit else ("", "")


def time_of_view(v: str, adj: bool):  # time.go:276-331; adj = the exclusive upper bound of the view's period
    """None for "" (the zero time). ValueError with the reference's text for a name that is not a time view."""
    if v == "":
        return None
    part = view_time_part(v)
    fmt = {4: "%Y", 6: "%Y%m", 8: "%Y%m%d", 10: "%Y%m%d%H"}.get(len(part))
    if fmt is None:
        raise ValueError(f"invalid time format on view: {v}")
    t = datetime.strptime(part, fmt)
    if adj:
        t = {4: lambda: _add_date(t, 1, 0, 0), 6: lambda: _add_month(t), 8: lambda: _add_date(t, 0, 0, 1), 10: lambda: t + timedelta(hours=1)}[len(part)]()
    return t
