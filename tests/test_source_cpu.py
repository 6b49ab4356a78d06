"""Columnar file sources (queryengine_b200/source.py), host side only: Arrow chunks become the buffers
qe_table_upload takes, and the reference's CSV dialect is parsed with its null / boolean / double rules
(M/data/UnivocityCsvTable.kt:53-64)."""
import numpy as np
import pyarrow as pa
import pytest

from queryengine_b200.ast import DataType
from queryengine_b200.engine import Field, StringDictionary
from queryengine_b200.source import chunk_buffers, field_of_arrow, read_reference_csv

D, B, S, I, T = DataType.DOUBLE, DataType.BOOLEAN, DataType.STRING, DataType.INT64, DataType.DATE32


def _bits(bitmap, n):
    return np.unpackbits(bitmap, bitorder="little")[:n].astype(bool)


def test_arrow_types_map_to_fields():
    assert field_of_arrow("a", pa.float64()).type == D
    assert field_of_arrow("a", pa.int32()).type == D            # cast to float64 on read
    assert field_of_arrow("a", pa.int64()).type == I
    assert field_of_arrow("a", pa.date32()).type == T
    assert field_of_arrow("a", pa.bool_()).type == B
    assert field_of_arrow("a", pa.string()).type == S
    assert field_of_arrow("a", pa.dictionary(pa.int8(), pa.string())).type == S
    with pytest.raises(TypeError):
        field_of_arrow("a", pa.list_(pa.int8()))


@pytest.mark.parametrize("offset", [0, 3, 8, 13])
def test_chunks_become_upload_buffers_at_any_slice_offset(offset):
    vals = [1.5, None, -0.0, 4.0, None, 6.25, 7.0, 8.0, 9.0, None, 11.0, 12.0, 13.0, 14.0, 15.0, 16.0, 17.0, None, 19.0]
    n = len(vals) - offset
    d = StringDictionary()
    a = pa.array(vals, pa.float64()).slice(offset)
    v, m = chunk_buffers(a, Field("x", D, True), d)
    want_valid = np.array([x is not None for x in vals[offset:]])
    assert np.array_equal(_bits(m, n), want_valid)
    assert all(v[i] == vals[offset + i] for i in range(n) if want_valid[i]) and v.dtype == np.float64
    # integers that are not int64 are cast to float64
    v, m = chunk_buffers(pa.array([1, 2, None, 4], pa.int16()), Field("x", D, True), d)
    assert v.dtype == np.float64 and list(v[[0, 1, 3]]) == [1.0, 2.0, 4.0] and list(_bits(m, 4)) == [True, True, False, True]
    # booleans: the value bitmap itself
    bools = [True, False, None, True, True, False, True, None, False, True, True][: max(1, 11 - offset % 8)]
    b = pa.array(bools, pa.bool_()).slice(offset % 3)
    v, m = chunk_buffers(b, Field("b", B, True), d)
    k = len(b)
    want = [x for x in bools[offset % 3:]]
    assert [bool(x) for x, ok in zip(_bits(v, k), _bits(m, k)) if ok] == [x for x in want if x is not None]
    # strings: codes of the context's dictionary, nulls in the bitmap
    strs = ["N", "R", None, "A", "R", "N", None, "zz", "A"]
    s = pa.array(strs, pa.string()).slice(offset % 4)
    v, m = chunk_buffers(s, Field("s", S, True), d)
    k = len(s)
    got = [d.decode(c) if ok else None for c, ok in zip(v, _bits(m, k))]
    assert got == strs[offset % 4:] and v.dtype == np.int32
    # the same strings arriving dictionary-encoded (Parquet / IPC dictionaries) give the same codes
    v2, m2 = chunk_buffers(s.dictionary_encode(), Field("s", S, True), d)
    assert np.array_equal(v2[_bits(m2, k)], v[_bits(m, k)])


def test_reference_csv_dialect(tmp_path):
    p = tmp_path / "t.csv"
    p.write_text("foo,bar,flag\n1.5,a,true\n,b,TRUE\n2e3,,false\n-0.0,d,yes\n4,e,\n")
    fields = [Field("foo", D, True), Field("bar", S, True), Field("flag", B, True)]
    t = read_reference_csv(str(p), fields)
    assert t.column("foo").to_pylist() == [1.5, None, 2000.0, -0.0, 4.0]
    assert t.column("bar").to_pylist() == ["a", "b", None, "d", "e"]
    assert t.column("flag").to_pylist() == [True, True, False, False, None]        # toBoolean(): anything but "true" is false
    t = read_reference_csv(str(p), fields, ["flag", "foo"])                        # settings.selectFields(projection)
    assert t.column_names == ["flag", "foo"]
    with pytest.raises(ValueError):
        read_reference_csv(str(p), fields, ["nope"])


def test_reference_csv_whitespace_and_java_numerals(tmp_path):
    """univocity trims values by default; toDouble() takes what java.lang.Double.parseDouble takes."""
    from queryengine_b200.source import java_parse_double
    p = tmp_path / "w.csv"
    p.write_text('foo,bar,flag\n 1.5 ,  a b ,  true \n1d,"",TRUE\n0x1p3, x,\tTrue\nNaN,   ,false\n-Infinity,q, yes \n+.5e1,r,\n')
    fields = [Field("foo", D, True), Field("bar", S, True), Field("flag", B, True)]
    t = read_reference_csv(str(p), fields)
    foo = t.column("foo").to_pylist()
    assert foo[:3] == [1.5, 1.0, 8.0] and foo[3] != foo[3] and foo[4:] == [float("-inf"), 5.0]
    assert t.column("bar").to_pylist() == ["a b", None, "x", None, "q", "r"]            # quoted empty and all-blank: null
    assert t.column("flag").to_pylist() == [True, True, True, False, False, None]
    for good, v in (("1e3", 1000.0), ("  -0.0\t", -0.0), ("0X.8P1", 1.0), ("3.F", 3.0), ("Infinity", float("inf"))):
        assert java_parse_double(good) == v
    for bad in ("inf", "nan", "1_000", "0x10", "", "1e", "--1", "1.5 x"):
        with pytest.raises(ValueError):
            java_parse_double(bad)
    q = tmp_path / "bad.csv"
    q.write_text("foo\n1.5\ninf\n")
    with pytest.raises(ValueError):
        read_reference_csv(str(q), [Field("foo", D, True)])
