"""Columnar file sources feeding the HBM-resident table.

The reference scans CSV files row by row, boxing every field
(src/main/java/net/jhorstmann/queryengine/operator/CsvSourceOperator.kt:48-80,
.../data/UnivocityCsvTable.kt:10-67).  Here a `Table` is backed by a columnar file -- Arrow IPC,
Parquet, or the reference's own CSV dialect parsed by pyarrow -- whose column chunks go to the GPU as
they are: Arrow's value buffers and LSB-first validity bitmaps are exactly the layout `qe_table_upload`
takes (include/qegpu.h), so a chunk is uploaded from the file's (memory-mapped) buffer without a
conversion pass; STRING columns are dictionary-encoded by pyarrow and re-coded into the context's
dictionary.  `getScanOperator(projection)` reads only the named columns (Table.getScanOperator,
.../data/Table.kt:8; column pruning as in MemoryTable.kt:8-18).

pyarrow does the file parsing on the host; the library only sees plain pointers.
"""
from __future__ import annotations

import os
from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np

from .ast import ColumnExpression, DataType
from .engine import ColumnarTable, Context, Field, GpuPipelineOperator, StringDictionary


def _arrow():
    import pyarrow as pa
    return pa


def field_of_arrow(name: str, typ, nullable: bool = True) -> Field:
    """Arrow type -> Field.  float64 -> DOUBLE, bool -> BOOLEAN, strings / dictionaries of strings -> STRING
    (the reference's three types, .../data/Schema.kt:3-5); int64 / date32 map to the INT64 / DATE32 extensions,
    other numeric types are cast to float64 when they are read."""
    pa = _arrow()
    if pa.types.is_boolean(typ):
        return Field(name, DataType.BOOLEAN, nullable)
    if pa.types.is_string(typ) or pa.types.is_large_string(typ) or \
            (pa.types.is_dictionary(typ) and (pa.types.is_string(typ.value_type) or pa.types.is_large_string(typ.value_type))):
        return Field(name, DataType.STRING, nullable)
    if pa.types.is_int64(typ):
        return Field(name, DataType.INT64, nullable)
    if pa.types.is_date32(typ):
        return Field(name, DataType.DATE32, nullable)
    if pa.types.is_floating(typ) or pa.types.is_integer(typ) or pa.types.is_decimal(typ):
        return Field(name, DataType.DOUBLE, nullable)
    raise TypeError(f"column {name}: Arrow type {typ} has no counterpart on the GPU path")


def _bitmap(buf, offset: int, n: int) -> Optional[np.ndarray]:
    """Arrow validity / boolean bitmap of n bits starting at bit `offset` -> byte-aligned uint8 array
    (a zero-copy view when the offset is a whole number of bytes)."""
    if buf is None:
        return None
    raw = np.frombuffer(buf, dtype=np.uint8)
    if offset % 8 == 0:
        return raw[offset // 8: offset // 8 + (n + 7) // 8]
    bits = np.unpackbits(raw[offset // 8: (offset + n + 7) // 8], bitorder="little")[offset % 8: offset % 8 + n]
    return np.packbits(bits, bitorder="little")


def chunk_buffers(chunk, field: Field, dictionary: Optional[StringDictionary]) -> Tuple[np.ndarray, Optional[np.ndarray]]:
    """One Arrow array -> (values, validity bitmap or None) in the layout of qe_table_upload: float64 / int64 /
    int32 arrays (views of the Arrow buffer where possible), BOOLEAN as an LSB-first bitmap, STRING as int32
    codes of `dictionary`."""
    pa = _arrow()
    n = len(chunk)
    valid = None
    if chunk.null_count:
        valid = _bitmap(chunk.buffers()[0], chunk.offset, n)
    t = field.type
    if t == DataType.STRING:
        if dictionary is None:
            raise ValueError("STRING columns need the context's StringDictionary")
        enc = chunk if pa.types.is_dictionary(chunk.type) else chunk.dictionary_encode()
        local = np.array([dictionary.encode(s) if s is not None else 0 for s in enc.dictionary.to_pylist()], dtype=np.int32)
        idx = enc.indices
        if idx.null_count:
            valid = _bitmap(idx.buffers()[0], idx.offset, n)
        ind = np.frombuffer(idx.buffers()[1], dtype=idx.type.to_pandas_dtype())[idx.offset: idx.offset + n]
        if idx.null_count:                      # index slots of nulls are undefined: keep them in range
            ind = np.where(np.unpackbits(valid, bitorder="little")[:n].astype(bool), ind, 0)
        codes = local[ind] if len(local) else np.zeros(n, np.int32)
        return np.ascontiguousarray(codes, dtype=np.int32), valid
    if t == DataType.BOOLEAN:
        return _bitmap(chunk.buffers()[1], chunk.offset, n), valid
    want = {DataType.DOUBLE: pa.float64(), DataType.INT64: pa.int64(), DataType.DATE32: pa.date32()}[t]
    if chunk.type != want:
        chunk = chunk.cast(want)
        if chunk.null_count:
            valid = _bitmap(chunk.buffers()[0], chunk.offset, n)
    np_t = {DataType.DOUBLE: np.float64, DataType.INT64: np.int64, DataType.DATE32: np.int32}[t]
    vals = np.frombuffer(chunk.buffers()[1], dtype=np_t)[chunk.offset: chunk.offset + n] if n else np.zeros(0, np_t)
    return vals, valid


def upload_arrow(ctx: Context, arrow_table, fields: Optional[Sequence[Field]] = None) -> ColumnarTable:
    """pyarrow.Table -> ColumnarTable in HBM, chunk by chunk (asynchronous uploads, one synchronize at the end)."""
    pa = _arrow()
    if fields is None:
        fields = [field_of_arrow(f.name, f.type, f.nullable) for f in arrow_table.schema]
    n = arrow_table.num_rows
    table = ColumnarTable(ctx, fields, n)
    keep = []                                   # host buffers must outlive the asynchronous copies
    for c, f in enumerate(fields):
        col = arrow_table.column(c)
        if col.null_count and not f.nullable:
            raise ValueError(f"column {f.name} is declared non-nullable but has nulls")
        # chunk boundaries that are not multiples of 8 rows would split a validity byte: one contiguous array then
        bitmap_involved = f.nullable or f.type == DataType.BOOLEAN
        sizes = [len(ch) for ch in col.chunks]
        if bitmap_involved and any(s % 8 for s in sizes[:-1]):
            col = pa.chunked_array([col.combine_chunks()]) if n else col
        row0 = 0
        for ch in col.chunks:
            if len(ch) == 0:
                continue
            vals, valid = chunk_buffers(ch, f, ctx.dictionary)
            if f.nullable and valid is None:
                valid = np.full((len(ch) + 7) // 8, 0xFF, np.uint8)
            keep += [vals, valid, ch]
            table.upload_raw(c, vals.ctypes.data, valid.ctypes.data if (valid is not None and f.nullable) else None, row0, len(ch))
            row0 += len(ch)
    ctx.synchronize()
    del keep
    return table


_JAVA_DECIMAL = r"^[+-]?(\d+\.?\d*|\.\d+)([eE][+-]?\d+)?$"
_JAVA_DOUBLE = None


def java_parse_double(text: str) -> float:
    """java.lang.Double.parseDouble / Kotlin String.toDouble(): optional sign, then `NaN`, `Infinity`, a decimal or a
    hexadecimal floating-point literal, an optional type suffix d / D / f / F; surrounding whitespace is trimmed.
    Anything else is a NumberFormatException (ValueError here) -- e.g. `inf`, `nan`, `1_000`, `0x10` (no exponent)."""
    import re
    global _JAVA_DOUBLE
    if _JAVA_DOUBLE is None:
        dec = r"(?:\d+\.?\d*|\.\d+)(?:[eE][+-]?\d+)?"
        hexa = r"0[xX](?:[0-9a-fA-F]+\.?[0-9a-fA-F]*|\.[0-9a-fA-F]+)[pP][+-]?\d+"
        _JAVA_DOUBLE = re.compile(r"^([+-]?)(NaN|Infinity|(" + hexa + r")|(" + dec + r"))[dDfF]?$")
    t = text.strip(" \t\n\r\x0b\x0c" + "".join(chr(c) for c in range(0, 32)))
    m = _JAVA_DOUBLE.match(t)
    if not m:
        raise ValueError(f'For input string: "{text}"')            # NumberFormatException
    sign, body, hexa, dec = m.group(1), m.group(2), m.group(3), m.group(4)
    if body == "NaN":
        return float("nan")
    if body == "Infinity":
        v = float("inf")
    elif hexa:
        v = float.fromhex(hexa)
    else:
        v = float(dec)
    return -v if sign == "-" else v


def read_reference_csv(path: str, fields: Sequence[Field], projection: Optional[Sequence[str]] = None):
    """The reference's CSV dialect (UnivocityCsvTable.kt:10-67 with univocity's default CsvParserSettings): comma
    separated, header row, leading and trailing whitespace of a value is dropped (ignoreLeading/TrailingWhitespaces
    are on by default), an empty field -- also a quoted empty one, `isNullOrEmpty` :55 -- is null, BOOLEAN is
    `toBoolean()` (true iff the text is "true" in any case, :59), DOUBLE is `toDouble()` (:60, the grammar of
    java.lang.Double.parseDouble: see java_parse_double; a malformed number raises like NumberFormatException).
    Returns a pyarrow.Table of the projected columns with the types of `fields`.
    Not reproduced: whitespace INSIDE quotes, which univocity keeps (pyarrow's reader does not say which values were
    quoted); such a value is trimmed here."""
    pa = _arrow()
    import pyarrow.compute as pc
    import pyarrow.csv as pcsv
    names = [f.name for f in fields]
    want = list(projection) if projection is not None else names
    for name in want:
        if name not in names:
            raise ValueError(f"Field {name} not found in schema")           # IllegalArgumentException, UnivocityCsvTable.kt:19
    conv = pcsv.ConvertOptions(column_types={n: pa.string() for n in want}, include_columns=want,
                               strings_can_be_null=True, null_values=[""], quoted_strings_can_be_null=True)
    raw = pcsv.read_csv(path, convert_options=conv)
    cols = []
    for name in want:
        f = fields[names.index(name)]
        c = pc.utf8_trim(raw.column(name), characters="".join(chr(x) for x in range(0, 33)))      # univocity: chars <= ' '
        c = pc.if_else(pc.equal(pc.utf8_length(c), 0), pa.scalar(None, pa.string()), c)           # all blanks -> empty -> null
        if f.type == DataType.DOUBLE:
            plain = pc.fill_null(pc.match_substring_regex(c, _JAVA_DECIMAL), True)
            if pc.all(plain).as_py() is not False:
                c = pc.cast(c, pa.float64())
            else:                                                           # NaN, Infinity, hex floats, d / f suffixes
                c = pa.array([None if v is None else java_parse_double(v) for v in c.to_pylist()], pa.float64())
        elif f.type == DataType.BOOLEAN:
            c = pc.equal(pc.utf8_lower(c), "true")
        cols.append(c)
    return pa.table(cols, names=want)


class FileTable:
    """Table { schema; getScanOperator(projection) } (.../data/Table.kt:6-10) over a columnar file.

    `path` ends in .arrow / .feather / .ipc (Arrow IPC file, memory-mapped), .parquet, or .csv (the reference's
    dialect; `fields` is then required, like the Schema a UnivocityCsvTable is constructed with)."""

    def __init__(self, ctx: Context, path: str, fields: Optional[Sequence[Field]] = None):
        self.ctx = ctx
        self.path = path
        self.kind = os.path.splitext(path)[1].lower().lstrip(".")
        if self.kind not in ("arrow", "feather", "ipc", "parquet", "csv"):
            raise ValueError(f"unknown file type: {path}")
        if self.kind == "csv" and fields is None:
            raise ValueError("a CSV table needs its schema")
        self._fields = list(fields) if fields is not None else None
        self._loaded = {}

    @property
    def fields(self) -> List[Field]:
        if self._fields is None:
            pa = _arrow()
            if self.kind == "parquet":
                import pyarrow.parquet as pq
                schema = pq.read_schema(self.path)
            else:
                with pa.memory_map(self.path) as src:
                    schema = pa.ipc.open_file(src).schema
            self._fields = [field_of_arrow(f.name, f.type, f.nullable) for f in schema]
        return self._fields

    def read(self, projection: Sequence[str]):
        """pyarrow.Table of the named columns only."""
        pa = _arrow()
        names = [f.name for f in self.fields]
        for name in projection:
            if name not in names:
                raise ValueError(f"Unknown field {name}")
        if self.kind == "csv":
            return read_reference_csv(self.path, self.fields, projection)
        if self.kind == "parquet":
            import pyarrow.parquet as pq
            return pq.read_table(self.path, columns=list(projection), memory_map=True)
        src = pa.memory_map(self.path)
        return pa.ipc.open_file(src).read_all().select(list(projection))

    def load(self, projection: Sequence[str]) -> ColumnarTable:
        """The named columns as an HBM-resident table (cached per projection)."""
        key = tuple(projection)
        if key not in self._loaded:
            names = [f.name for f in self.fields]
            fields = [self.fields[names.index(n)] for n in projection]
            self._loaded[key] = upload_arrow(self.ctx, self.read(projection), fields)
        return self._loaded[key]

    def getScanOperator(self, projection: Sequence[str]) -> GpuPipelineOperator:
        t = self.load(projection)
        exprs = [ColumnExpression(f.name, i, f.type) for i, f in enumerate(t.fields)]
        return GpuPipelineOperator(t, None, exprs, None)

    def close(self) -> None:
        for t in self._loaded.values():
            t.close()
        self._loaded = {}
