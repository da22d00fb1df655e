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


def write_table(frame, path, file_type: str = "csv"):
    """write_df_to_file, src/utils.rs:121-144: csv / parquet / json / ndjson."""
    import pyarrow as pa

    at = frame if isinstance(frame, pa.Table) else frame.to_arrow()
    if file_type == "csv":
        # the reference writes through polars' CsvWriter (utils.rs:121-144): fields quoted only when they need it
        try:
            import polars as pl

            (frame if not isinstance(frame, pa.Table) else pl.from_arrow(frame)).write_csv(path)
        except ImportError:
            import pyarrow.csv as pacsv

            pacsv.write_csv(at, path, write_options=pacsv.WriteOptions(quoting_style="needed"))
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
