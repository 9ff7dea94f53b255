"""CPU restatement of the reference's Aggregate WITHOUT group-by columns, per aggregate column (numpy). TEST
INFRASTRUCTURE: imported only by tests/ — never by the product package.

Follows src/lib/operators/aggregate.cpp: _aggregate_column (:203-252: rows are visited in chunk order, a NULL row leaves
the aggregate untouched, every other row goes through the functor and bumps aggregate_count), the functors (:141-201:
MIN / MAX replace on value_smaller / value_greater, SUM adds, AVG reuses SUM), the result writers (:622-712: MIN / MAX /
SUM / AVG are NULL when no row contributed, AVG = aggregate / count, COUNT = aggregate_count) and the accumulator types of
operators/aggregate/aggregate_traits.hpp:16-66 (COUNT int64; MIN / MAX the column type; SUM int64 for integers, double
for floats; AVG double for every arithmetic type). COUNT(*) is the row count (:424-431)."""
import numpy as np


def aggregate_column(values, is_null=None):
    """values: 1-D numpy array of the column's type (int32 / int64 / float32 / float64), is_null: bool / uint8 array or
    None. Returns {"count_rows", "count", "sum", "min", "max", "avg"} with None for SQL NULL."""
    values = np.asarray(values)
    keep = np.ones(len(values), bool) if is_null is None else ~np.asarray(is_null).astype(bool)
    v = values[keep]
    out = {"count_rows": int(len(values)), "count": int(len(v)), "sum": None, "min": None, "max": None, "avg": None}
    if len(v) == 0:
        return out
    if values.dtype.kind == "f":
        acc = np.add.accumulate(v.astype(np.float64))[-1]   # sequential double accumulation in row order (:173-175)
        out["sum"] = float(acc)
        out["avg"] = float(acc) / len(v)
        if np.isnan(v).any():
            lo = hi = v[0]
            for x in v[1:]:   # plain < and > on the column type: a NaN replaces nothing and is replaced by nothing
                if x < lo:
                    lo = x
                if x > hi:
                    hi = x
        else:                 # without NaNs the row-by-row replacement is the minimum / maximum
            lo, hi = v.min(), v.max()
        out["min"], out["max"] = float(lo), float(hi)
    else:
        total = int(v.astype(np.int64).sum(dtype=np.int64))   # int64_t accumulator, wraps on overflow
        out["sum"] = total
        out["avg"] = float(np.add.accumulate(v.astype(np.float64))[-1]) / len(v)   # AVG accumulates in double
        out["min"], out["max"] = int(v.min()), int(v.max())
    return out


def aggregate_grouped(columns, group_by, aggregates):
    """Aggregate WITH group-by columns (aggregate.cpp:254-420: the group key of a row is the tuple of its group-by values,
    NULL being a value of its own; groups appear in first-occurrence order, which the reference does not guarantee either).
    columns: {name: (values, is_null)}; group_by: [name]; aggregates: [(function, name or None for COUNT(*))] with function in
    "min" / "max" / "sum" / "avg" / "count". Returns a list of rows: group-by values (None = NULL) followed by the aggregates,
    with the types and NULL rules of aggregate_column()."""
    n_rows = len(next(iter(columns.values()))[0]) if columns else 0
    groups, order = {}, []
    for i in range(n_rows):
        key = tuple(None if columns[g][1][i] else columns[g][0][i].item() for g in group_by)
        if key not in groups:
            groups[key] = []
            order.append(key)
        groups[key].append(i)
    rows = []
    for key in order:
        idx = np.array(groups[key], dtype=np.int64)
        row = list(key)
        for function, name in aggregates:
            if name is None:                       # COUNT(*): rows of the group (aggregate.cpp:424-431)
                row.append(len(idx))
                continue
            values, is_null = columns[name]
            result = aggregate_column(values[idx], np.asarray(is_null)[idx])
            row.append(result[function])
        rows.append(row)
    return rows
