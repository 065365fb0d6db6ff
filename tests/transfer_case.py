"""Shared by tests/test_transfer.py (host loops over the per-item routines, no GPU) and
tests/test_transfer_gpu.py (the C-ABI on the device): run a recorded case of the reference's remesh
field transfer (tests/transfer_capture.py) through `cells(...)` / `nodes(...)` and compare."""
import os
import subprocess
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(HERE, "golden", "transfer-t3d-remesh-coarse.npz")
REFLIB = os.path.join(ROOT, "oracle", "_ref", "libdesref.so")


def load_case(path):
    d = np.load(path)
    c = {k: d[k] for k in d.files}
    # the slim fixture keeps one copy of what the three calls share
    for pre in ("nn1_", "brc_"):
        for nm in ("old_coord", "new_coord"):
            c.setdefault(pre + nm, c["nn0_" + nm])
    c.setdefault("brc_old_conn", c["nn0_old_conn"])
    c.setdefault("brc_old_size", c["nn0_old_size"])
    return c


def live_case(cfg, steps, tmp_path, extra=()):
    """The same recording made now by the reference compiled under oracle/_ref (None if absent).
    cfg: a file under tests/cfg or an absolute path; extra: more arguments for transfer_capture.py."""
    if not os.path.exists(REFLIB):
        return None
    out = os.path.join(str(tmp_path), "case.npz")
    subprocess.run([sys.executable, os.path.join(HERE, "transfer_capture.py"), os.path.join(HERE, "cfg", cfg),
                    str(steps), out] + list(extra), check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=900,
                   env=dict(os.environ, OMP_NUM_THREADS=str(min(8, os.cpu_count() or 1))))
    return load_case(out)


def fields_of(c, pre):
    return {k[len(pre) + 4:]: c[k] for k in c if k.startswith(pre + "src_")}


def check_cells(c, pre, is_surface, cells):
    """cells(old_coord, old_conn, old_size, new_coord, new_conn, fields, is_surface) ->
    (outputs, is_changed, stats).  Everything must equal the reference's bit for bit."""
    flds = fields_of(c, pre)
    assert flds
    out, changed, st = cells(c[pre + "old_coord"], c[pre + "old_conn"], c[pre + "old_size"],
                             c[pre + "new_coord"], c[pre + "new_conn"], flds, is_surface)
    if pre + "is_changed" in c:
        assert np.array_equal(changed, c[pre + "is_changed"]), "is_changed"
    for name in flds:
        ref = c[pre + "dst_" + name]
        assert out[name].shape == ref.shape, name
        assert np.array_equal(out[name], ref), (pre, name, int((out[name] != ref).sum()))
    assert st.nchanged == int((changed != 0).sum())
    return changed, st


def check_nodes(c, pre, nodes):
    """nodes(old_coord, old_conn, old_size, support_idx, support_arr, new_coord, fields) ->
    (outputs, el, brc, stats)."""
    flds = fields_of(c, pre)
    assert flds
    out, el, brc, st = nodes(c[pre + "old_coord"], c[pre + "old_conn"], c[pre + "old_size"],
                             c[pre + "old_support_idx"], c[pre + "old_support_arr"], c[pre + "new_coord"], flds)
    for name in flds:
        ref = c[pre + "dst_" + name]
        assert np.array_equal(out[name], ref), (pre, name, int((out[name] != ref).sum()))
    nold = c[pre + "old_conn"].size // 4
    assert el.min() >= 0 and el.max() < nold
    assert np.all(np.abs(brc.sum(axis=1) - 1) < 1e-12)
    return el, brc, st


def tiny_box_case(resolution, tmp_path):
    """box-ep-tiny (12 x 8 x 6 km) at `resolution`, transferred onto its jittered self: at 6 km the
    old mesh has 10 elements, fewer than the 32 neighbours find_acm_elem_ratios asks for
    (max_el = min(32, old_nelem), nn-interpolation.cxx:99)."""
    from conftest import cfg_variant
    cfg, _ = cfg_variant("box-ep-tiny", {"mesh.resolution": resolution})
    return live_case(cfg, 3, tmp_path, extra=("--jitter", "0.15"))
