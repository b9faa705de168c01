"""libgeocb.so loads without a GPU and exports exactly the symbols include/geocb.h declares."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def header_functions():
    txt = (ROOT / "include" / "geocb.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(geocb_[a-z0-9_]+)\s*\(", txt)))


def test_header_and_binding_agree():
    from vqvae_b200 import _lib
    assert header_functions() == sorted(_lib.SIGNATURES)


def test_library_exports_every_symbol():
    from vqvae_b200 import _lib
    lib = _lib.load()
    for name in header_functions():
        assert hasattr(lib, name), name
    assert lib.geocb_abi_version() == 1


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from vqvae_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()


def test_product_never_imports_oracle():
    for p in (ROOT / "vqvae_b200").rglob("*.py"):
        s = p.read_text()
        assert "geo_oracle" not in s and "import oracle" not in s and "from oracle" not in s, p
        assert not re.search(r"^\s*(import|from)\s+(sklearn|scipy\.sparse\.csgraph)", s, flags=re.M), p


def np_pairwise_tree(x, plan):
    """Evaluate the C plan with NumPy float32 arithmetic."""
    nl, nn, nlev = int(plan[1]), int(plan[2]), int(plan[3])
    leaves = plan[5:5 + 2 * nl].reshape(nl, 2)
    lvl = plan[5 + 2 * nl: 5 + 2 * nl + nlev + 1]
    nodes = plan[5 + 2 * nl + nlev + 1:][: 2 * nn].reshape(nn, 2)
    vals = np.zeros(nl + nn, np.float32)
    for i, (off, ln) in enumerate(leaves):
        a = x[off:off + ln]
        if ln < 8:
            r = np.float32(0)
            for v in a:
                r = np.float32(r + v)
        else:
            body = ln - ln % 8
            acc = a[:8].copy()
            for j in range(8, body, 8):
                acc = (acc + a[j:j + 8]).astype(np.float32)
            r = np.float32(np.float32(np.float32(acc[0] + acc[1]) + np.float32(acc[2] + acc[3]))
                           + np.float32(np.float32(acc[4] + acc[5]) + np.float32(acc[6] + acc[7])))
            for v in a[body:]:
                r = np.float32(r + v)
        vals[i] = r
    for k in range(nlev):
        for pos in range(lvl[k], lvl[k + 1]):
            vals[nl + pos] = np.float32(vals[nodes[pos, 0]] + vals[nodes[pos, 1]])
    return vals[int(plan[4])]


@pytest.mark.parametrize("n", [1, 5, 8, 127, 128, 129, 1000, 4097, 20000, 123457])
def test_pairwise_plan_matches_numpy_sum(n):
    """The host-built reduction tree (geocb_pairwise_plan_host) is NumPy's float32 pairwise sum."""
    from vqvae_b200 import _lib
    lib = _lib.load()
    words = lib.geocb_pairwise_plan_size(n)
    buf = (ctypes.c_int64 * words)()
    assert lib.geocb_pairwise_plan_host(n, buf) == 0
    plan = np.frombuffer(buf, dtype=np.int64)
    x = (np.random.RandomState(n).rand(n).astype(np.float32) ** 2) * 7
    assert np_pairwise_tree(x, plan) == np.sum(x)
