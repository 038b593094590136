"""Shared loader for tests/golden/reference_kat.json (the reference's known-answer tests)."""
import json
import os

import numpy as np

from oracle.oracle import HostColumn

HERE = os.path.dirname(os.path.abspath(__file__))


def _arr(vals, dtype):
    dt = np.dtype(dtype)
    out = np.zeros(len(vals), dtype=dt)
    for i, v in enumerate(vals):
        if isinstance(v, str):
            v = float(v)
        if dt.kind == "b":
            out[i] = bool(v)
        elif dt.kind in "iu":
            out[i] = np.array(int(v)).astype(dt) if not isinstance(v, float) else int(v)
        else:
            out[i] = v
    return out


def load_cases(filename="reference_kat.json"):
    with open(os.path.join(HERE, "golden", filename)) as f:
        doc = json.load(f)
    out = []
    for c in doc["cases"]:
        for kt in c["key_types"]:
            for vt in c["value_types"]:
                if c["name"].startswith("m2_") and len(c["key_types"]) == 1 and vt not in ("float64",) and "nan" not in str(c["vals"]):
                    pass
                out.append((f'{c["name"]}[{kt}-{vt}]', c, kt, vt))
    return out


def materialize(c, kt, vt):
    """-> keys HostColumn, values HostColumn (honouring offset/length slices), expectation dict."""
    off, ln = c["offset"], c["length"]
    kdata = _arr(c["keys"], kt)
    vdata = _arr(c["vals"], vt)
    kvalid = None if c["key_valid"] is None else np.array(c["key_valid"], dtype=bool)
    vvalid = None if c["val_valid"] is None else np.array(c["val_valid"], dtype=bool)
    if ln is not None:  # cudf::slice: same buffers, offset/size view -> we slice the host arrays
        kdata, vdata = kdata[off:off + ln], vdata[off:off + ln]
        kvalid = None if kvalid is None else kvalid[off:off + ln]
        vvalid = None if vvalid is None else vvalid[off:off + ln]
    keys = HostColumn(kdata, kvalid)
    vals = HostColumn(vdata, vvalid)
    exp_keys = _arr(c["expect_keys"], kt)
    exp_kvalid = None if c["expect_key_valid"] is None else np.array(c["expect_key_valid"], dtype=bool)
    exp_vals = np.array([float(x) if isinstance(x, str) else x for x in c["expect"]], dtype=np.float64) \
        if any(isinstance(x, (str, float)) for x in c["expect"]) else np.array(c["expect"], dtype=np.int64)
    exp_valid = None if c["expect_valid"] is None else np.array(c["expect_valid"], dtype=bool)
    return keys, vals, dict(keys=exp_keys, key_valid=exp_kvalid, vals=exp_vals, valid=exp_valid, rtol=c["rtol"])


def check(name, G, kout, aout, exp):
    """Compare a key-sorted single-key / single-agg result with the expectation."""
    kd, kv, _ = kout
    ad, av, _ = aout
    assert G == len(exp["keys"]), f"{name}: {G} groups, expected {len(exp['keys'])}"
    ekv = exp["key_valid"] if exp["key_valid"] is not None else np.ones(G, dtype=bool)
    np.testing.assert_array_equal(kv, ekv, err_msg=f"{name}: key validity")
    np.testing.assert_array_equal(kd[ekv], exp["keys"][ekv], err_msg=f"{name}: keys")
    ev = exp["valid"] if exp["valid"] is not None else np.ones(G, dtype=bool)
    np.testing.assert_array_equal(av, ev, err_msg=f"{name}: result validity")
    got = ad[ev].astype(np.float64) if ad.dtype.kind == "f" or exp["vals"].dtype.kind == "f" else ad[ev].astype(np.int64)
    want = exp["vals"][ev]
    if ad.dtype.kind == "b":
        want = want.astype(bool)
    if exp["rtol"] > 0 or got.dtype.kind == "f":
        rt = max(exp["rtol"], 1e-6 if ad.dtype == np.float32 else 0.0)
        np.testing.assert_allclose(got, want.astype(np.float64), rtol=rt, atol=0, equal_nan=True, err_msg=name)
    else:
        np.testing.assert_array_equal(got, want, err_msg=name)
