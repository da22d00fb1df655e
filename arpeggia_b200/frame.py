"""Result table -> Arrow / polars with the reference's schema (src/contacts/mod.rs:140-214)."""
from __future__ import annotations

import numpy as np

from .core import COLUMNS, STRING_COLS


def to_arrow(table: dict):
    """dict of numpy columns (core.read_table) -> pyarrow.Table: model u32, distance/sc_* f32, resi/atomi i32,
    string columns large_utf8 (polars' String), sc_* null where either residue has no side-chain plane."""
    import pyarrow as pa

    cols, names = [], []
    for name in COLUMNS:
        v = table[name]
        if name == "model":
            a = pa.array(v, type=pa.uint32())
        elif name == "interaction" or name.split("_", 1)[-1] in STRING_COLS:
            a = pa.array(v.tolist(), type=pa.large_string())
        elif name.startswith("sc_"):
            a = pa.array(v, type=pa.float32(), mask=~table["sc_valid"])
        elif name == "distance":
            a = pa.array(v, type=pa.float32())
        else:
            a = pa.array(v, type=pa.int32())
        cols.append(a)
        names.append(name)
    return pa.table(cols, names=names)


def to_frame(table: dict):
    """polars.DataFrame when polars is importable (the reference's return type), else the pyarrow.Table."""
    at = to_arrow(table)
    try:
        import polars as pl
    except ImportError:
        return at
    return pl.from_arrow(at)


def csv_text(at) -> str:
    """The table as CSV text in the layout of polars' CsvWriter (what the reference's CLI writes by default): built column by
    column in Arrow's C++ (pyarrow's own CSV writer quotes every string and prints 2.0 as 2)."""
    import pyarrow as pa
    import pyarrow.compute as pc

    n = at.num_rows
    if n == 0:
        return ",".join(at.column_names) + "\n"
    big = pa.large_string()
    cols = []
    for name in at.column_names:
        col = at[name].combine_chunks()
        if pa.types.is_dictionary(col.type):
            col = col.dictionary_decode()
        if pa.types.is_floating(col.type):
            s = pc.cast(col, big)
            whole = pc.match_substring_regex(s, r"^-?[0-9]+$")  # an integral value: Arrow prints no fraction, polars does
            s = pc.if_else(whole, pc.binary_join_element_wise(s, pa.scalar(".0", big), pa.scalar("", big)), s)
            s = pc.replace_substring_regex(s, r"e\+?(-?)0*([0-9])", r"e\1\2")  # 1e-07 -> 1e-7, 1e+16 -> 1e16
        elif pa.types.is_string(col.type) or pa.types.is_large_string(col.type):
            s = pc.cast(col, big)
            needs = pc.match_substring_regex(s, '[",\n\r]')
            if pc.any(needs).as_py():
                q = pa.scalar('"', big)
                quoted = pc.binary_join_element_wise(q, pc.replace_substring(s, '"', '""'), q, pa.scalar("", big))
                s = pc.if_else(needs, quoted, s)
        else:
            s = pc.cast(col, big)
        cols.append(pc.fill_null(s, ""))
    rows = pc.binary_join_element_wise(*cols, pa.scalar(",", big))
    body = pc.binary_join(pa.LargeListArray.from_arrays(pa.array([0, n], type=pa.int64()), rows), pa.scalar("\n", big))[0].as_py()
    return ",".join(at.column_names) + "\n" + body + "\n"


def write_table(frame, path, file_type: str = "csv"):
    """write_df_to_file, src/utils.rs:121-144: csv / parquet / json / ndjson."""
    import pyarrow as pa

    at = frame if isinstance(frame, pa.Table) else frame.to_arrow()
    if file_type == "csv":
        # the reference writes through polars' CsvWriter (utils.rs:121-144): bare header, fields quoted only when they need
        # it, floats in their shortest form with a fraction ("2.0"), nulls as empty fields
        try:
            import polars as pl

            (frame if not isinstance(frame, pa.Table) else pl.from_arrow(frame)).write_csv(path)
        except ImportError:
            with open(path, "w", newline="") as f:
                f.write(csv_text(at))
    elif file_type == "parquet":
        import pyarrow.parquet as pq

        pq.write_table(at, path)
    elif file_type in ("json", "ndjson"):
        import json

        rows = at.to_pylist()
        with open(path, "w") as f:
            if file_type == "json":
                json.dump(rows, f)
            else:
                for r in rows:
                    f.write(json.dumps(r) + "\n")
    else:
        raise ValueError(f"unknown output format {file_type!r}")
