"""Random Arrow arrays for the differential tests against the compiled reference (tests/test_ref_scan.py, test_ref_append.py):
types to depth 3 — lists, large lists, structs, maps, fixed-size lists, optionally dictionaries — over 26 leaf types, random NULL
rates.  Dictionaries are optional because two shapes with them make the REFERENCE overflow a buffer."""
import decimal

import numpy as np
import pyarrow as pa

ALLOW_DICTIONARY = False


def rand_type(rng, depth=0):
    leaf = ["i8","i16","i32","i64","u8","u16","u32","u64","f32","f64","bool","str","lstr","bin","fsb","date32","date64","ts_us","ts_ms","ts_ns","time64","dec9","dec18","dec38","dur","sview"]
    nested = ["list","llist","struct","map","fsl"] + (["dict"] if ALLOW_DICTIONARY else [])
    if depth < 2 and rng.random() < 0.45:
        return (nested[int(rng.integers(0,len(nested)))], depth)
    return (leaf[int(rng.integers(0,len(leaf)))], depth)

def gen(rng, n, depth=0, allow_null=True):
    kind,_ = rand_type(rng, depth)
    p_null = float(rng.choice([0.0, 0.1, 0.5])) if allow_null else 0.0
    mask = rng.random(n) < p_null
    def m(vals, ty): return pa.array(vals, ty, mask=mask) if p_null>0 else pa.array(vals, ty)
    if kind in ("i8","i16","i32","i64","u8","u16","u32","u64"):
        dt = {"i8":np.int8,"i16":np.int16,"i32":np.int32,"i64":np.int64,"u8":np.uint8,"u16":np.uint16,"u32":np.uint32,"u64":np.uint64}[kind]
        info=np.iinfo(dt); return m(rng.integers(info.min, info.max, n, dtype=dt, endpoint=True), None)
    if kind=="f32": return m(rng.standard_normal(n).astype(np.float32), None)
    if kind=="f64": return m(rng.standard_normal(n), None)
    if kind=="bool": return m(rng.random(n)<0.5, None)
    if kind in ("str","lstr","sview"):
        ty={"str":pa.string(),"lstr":pa.large_string(),"sview":pa.string_view()}[kind]
        vals=[None if mask[i] else "x"*int(k) + str(i) for i,k in enumerate(rng.integers(0,25,n))]
        return pa.array(vals, ty)
    if kind=="bin": return pa.array([None if mask[i] else bytes(rng.integers(0,256,int(k)).astype(np.uint8)) for i,k in enumerate(rng.integers(0,20,n))], pa.binary())
    if kind=="fsb": return pa.array([None if mask[i] else bytes(rng.integers(0,256,3).astype(np.uint8)) for i in range(n)], pa.binary(3))
    if kind=="date32": return m(rng.integers(-100000,100000,n).astype(np.int32), pa.int32()).view(pa.date32())
    if kind=="date64": return m(rng.integers(-10000,10000,n)*86400000, pa.int64()).view(pa.date64())
    if kind=="ts_us": return m(rng.integers(-2**50,2**50,n), pa.int64()).view(pa.timestamp("us"))
    if kind=="ts_ms": return m(rng.integers(-2**40,2**40,n), pa.int64()).view(pa.timestamp("ms"))
    if kind=="ts_ns": return m(rng.integers(-2**60,2**60,n), pa.int64()).view(pa.timestamp("ns"))
    if kind=="time64": return m(rng.integers(0,86400000000,n), pa.int64()).view(pa.time64("us"))
    if kind=="dur": return m(rng.integers(-2**40,2**40,n), pa.int64()).view(pa.duration("ms"))
    if kind.startswith("dec"):
        w={"dec9":(9,2),"dec18":(18,4),"dec38":(38,6)}[kind]
        D=decimal.Decimal
        with decimal.localcontext() as c:
            c.prec=60
            vals=[None if mask[i] else D(int(x)).scaleb(-w[1]) for i,x in enumerate(rng.integers(-10**(min(w[0],18)-1),10**(min(w[0],18)-1),n))]
        return pa.array(vals, pa.decimal128(*w))
    if kind in ("list","llist"):
        lens=rng.integers(0,4,n); offs=np.concatenate([[0],np.cumsum(lens)])
        child=gen(rng,int(offs[-1]),depth+1)
        cls=pa.ListArray if kind=="list" else pa.LargeListArray
        o=pa.array(offs.astype(np.int32 if kind=="list" else np.int64))
        arr=cls.from_arrays(o, child)
        if p_null>0:
            arr=cls.from_arrays(pa.array([None if (i<n and mask[i]) else int(v) for i,v in enumerate(offs)], o.type), child)
        return arr
    if kind=="struct":
        k=int(rng.integers(1,4))
        kids=[gen(rng,n,depth+1) for _ in range(k)]
        return pa.StructArray.from_arrays(kids, [f"f{i}" for i in range(k)], mask=pa.array(mask) if p_null>0 else None)
    if kind=="map":
        lens=rng.integers(0,3,n); offs=np.concatenate([[0],np.cumsum(lens)]).astype(np.int32)
        keys=pa.array([f"k{i}" for i in range(int(offs[-1]))])
        vals=gen(rng,int(offs[-1]),depth+1)
        return pa.MapArray.from_arrays(pa.array(offs), keys, vals)
    if kind=="fsl":
        k=int(rng.integers(1,4)); child=gen(rng,n*k,depth+1)
        return pa.FixedSizeListArray.from_arrays(child, k, mask=pa.array(mask) if p_null>0 else None)
    if kind=="dict":
        dv=gen(rng,int(rng.integers(1,6)),2, allow_null=False)
        idx=pa.array(rng.integers(0,len(dv),n).astype(np.int16), mask=mask) if p_null>0 else pa.array(rng.integers(0,len(dv),n).astype(np.int16))
        return pa.DictionaryArray.from_arrays(idx, dv)
    raise AssertionError(kind)



# ---- random Arrow TYPES over the whole zoo the type mapping knows (and a few it must refuse), for the schema-level comparison
LEAVES=[pa.int8(),pa.int16(),pa.int32(),pa.int64(),pa.uint8(),pa.uint16(),pa.uint32(),pa.uint64(),pa.float16(),pa.float32(),pa.float64(),
 pa.bool_(),pa.string(),pa.large_string(),pa.string_view(),pa.binary(),pa.large_binary(),pa.binary_view(),pa.binary(7),pa.date32(),pa.date64(),
 pa.time32("s"),pa.time32("ms"),pa.time64("us"),pa.time64("ns"),pa.timestamp("s"),pa.timestamp("ms"),pa.timestamp("us"),pa.timestamp("ns"),
 pa.timestamp("us",tz="UTC"),pa.timestamp("ns",tz="Europe/Amsterdam"),pa.duration("s"),pa.duration("ms"),pa.duration("us"),pa.duration("ns"),
 pa.month_day_nano_interval(),pa.decimal128(4,1),pa.decimal128(9,2),pa.decimal128(18,3),pa.decimal128(38,10),pa.decimal256(40,3),
 pa.decimal32(7,2),pa.decimal64(15,2),pa.null(),pa.uuid(),pa.bool8(),pa.json_(),pa.opaque(pa.binary(16),"hugeint","DuckDB"),
 pa.opaque(pa.binary(16),"uhugeint","DuckDB"),pa.opaque(pa.binary(8),"time_tz","DuckDB"),pa.opaque(pa.binary(),"bit","DuckDB"),
 pa.opaque(pa.binary(),"varint","DuckDB"),pa.opaque(pa.int32(),"other","Vendor")]
def random_type(rng, depth=0):
    if depth<3 and rng.random()<0.45:
        k=int(rng.integers(0,10))
        if k==0: return pa.list_(random_type(rng, depth + 1))
        if k==1: return pa.large_list(random_type(rng, depth + 1))
        if k==2: return pa.struct([pa.field(nm, random_type(rng, depth + 1)) for nm in ["a","B b","c"][:int(rng.integers(1,4))]])
        if k==3: return pa.map_(pa.string() if rng.random()<0.7 else pa.int32(), random_type(rng, depth + 1))
        if k==4: return pa.list_(random_type(rng, depth + 1), int(rng.integers(1,4)))
        if k==5: return pa.dictionary(pa.int8() if rng.random()<0.5 else pa.uint32(), LEAVES[int(rng.integers(0,30))])
        if k==6: return pa.list_view(random_type(rng, depth + 1))
        if k==7: return pa.large_list_view(random_type(rng, depth + 1))
        if k==8: return pa.run_end_encoded(pa.int32(), LEAVES[int(rng.integers(0,30))])
        if k==9: return pa.union([pa.field("x", random_type(rng, depth + 1)), pa.field("y", LEAVES[int(rng.integers(0,30))])], mode="sparse")
    return LEAVES[int(rng.integers(0,len(LEAVES)))]
def empty_array(t):
    if isinstance(t, pa.BaseExtensionType):
        return pa.ExtensionArray.from_storage(t, empty_array(t.storage_type))
    return pa.array([], t)
