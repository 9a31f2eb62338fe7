"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports exactly the
symbols include/piccolo_b200.h declares, and refuses to compute without a device (no CPU fallback)."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def built():
    from piccolo_b200 import build as B
    B.build_cuda()
    return B.LIB


def _declared():
    hdr = open(os.path.join(ROOT, "include", "piccolo_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(pcl_[a-z0-9_]+)\s*\(", hdr)))


def test_header_symbols_are_exported_and_bound(built):
    from piccolo_b200 import capi
    L = capi.lib()
    declared = _declared()
    assert len(declared) >= 30
    exported = subprocess.check_output(["nm", "-D", "--defined-only", built], text=True)
    exported = set(re.findall(r" T (pcl_[a-z0-9_]+)", exported))
    assert set(declared) <= exported, "declared but not exported: %s" % sorted(set(declared) - exported)
    assert set(declared) == set(capi.SIGNATURES), "binding table out of sync with the header"
    for name in declared:
        assert getattr(L, name) is not None
    assert L.pcl_version() >= 100


def test_built_for_sm100a_with_tma(built):
    """The cubin is sm_100a and the LBS kernel really uses the TMA bulk-copy path (SASS: UBLKCP)."""
    out = subprocess.run(["cuobjdump", "-sass", built], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout
    assert "UBLKCP" in out.stdout
    assert "lbs_kernel" in out.stdout and "pose_kernel" in out.stdout


def test_no_cpu_fallback(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from piccolo_b200 import capi
    with pytest.raises(capi.PclError) as e:
        capi.Context(0)
    assert e.value.code == -5 and "no CPU path" in str(e.value)


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under piccolo_b200/ may import, link or open it."""
    pkg = os.path.join(ROOT, "piccolo_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "liboracle" not in src and "pyoracle" not in src and "libpiccolo_ref" not in src, os.path.join(dp, f)


def test_synthetic_assets_are_deterministic_and_well_formed(player):
    from piccolo_b200 import assets as A
    sk, cl, am, mk = player
    assert sk.J == 58 and cl.C == 58 and cl.total_frame == 23 and sk.is_flat and sk.in_topological_order
    assert [k.shape[0] for k in cl.rot_keys].count(23) >= 57
    a, b = A.make_variant_clip(cl, 17, seed=1234), A.make_variant_clip(cl, 17, seed=1234)
    assert all(np.array_equal(x, y) for x, y in zip(a.rot_keys, b.rot_keys)) and a.total_frame == 17
    m = A.make_player_mesh(sk, V=4095, seed=42)
    assert m.V == 4095 and m.bind_idx.min() >= 0 and m.bind_idx.max() <= 58 and np.all(m.bind_idx[:, 0] > 0)
    assert np.array_equal(m.index_buffer, np.arange(4095))  # fully expanded, render_resource.cpp:705-712
    s5 = A.make_synthetic_skeleton(256)
    assert np.all(s5.parent_index[1:] < np.arange(1, 256)) and np.all(s5.tpose[:, 12:] == [0, 0, 0, 1])
