"""CPU-only: the C-ABI shared library loads and exports every symbol include/monogs_b200.h declares, and the
ctypes signature table agrees with the header (no compute calls — there is no GPU here)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "monogs_b200.h")


def _declared():
    """{name: n_args} for every function prototype in the header."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"#.*", "", src)
    out = {}
    for m in re.finditer(r"\b(?:int|size_t|const char\*)\s+(mgs_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    return out


def _lib():
    from monogs_b200 import _lib as L
    if not os.path.exists(L.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return L


def test_header_parses_and_is_nonempty():
    d = _declared()
    assert len(d) >= 18 and "mgs_raster_preprocess" in d and "mgs_raster_backward_pose" in d


def test_library_exports_every_declared_symbol():
    L = _lib()
    lib = C.CDLL(L.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in the header but not exported"


def test_ctypes_table_matches_header():
    L = _lib()
    d = _declared()
    assert set(L.SIGNATURES) == set(d), set(L.SIGNATURES) ^ set(d)
    for name, n in d.items():
        assert len(L.SIGNATURES[name][1]) == n, f"{name}: header has {n} args, ctypes table {len(L.SIGNATURES[name][1])}"


def test_struct_layouts_match_header_field_order():
    L = _lib()
    src = open(HEADER).read()
    for cname, cls in (("mgs_view", L.MgsView), ("mgs_loss_args", L.MgsLossArgs), ("mgs_adam_group", L.MgsAdamGroup),
                       ("mgs_gaussians", L.MgsGaussians), ("mgs_map_args", L.MgsMapArgs), ("mgs_map_view", L.MgsMapView)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), src, flags=re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for stmt in body.split(";"):
            stmt = stmt.strip()
            if not stmt:
                continue
            for part in stmt.split(","):
                names.append(re.findall(r"(\w+)\s*$", part.strip())[0])
        assert names == [f[0] for f in cls._fields_], (cname, names)


def test_error_path_without_gpu():
    """Argument validation happens before any CUDA call, and failures carry a message."""
    L = _lib()
    lib = L.lib()
    assert lib.mgs_version() >= 100
    rc = lib.mgs_raster_preprocess(None, 1, None, None, None, None, None, None, None, None, None, 0, None, None)
    assert rc == -1 and b"view" in lib.mgs_last_error_string()
    with pytest.raises(RuntimeError, match="view is null"):
        L.check(rc, "mgs_raster_preprocess")
    assert lib.mgs_adam_step(None, 0, 1, 0.9, 0.999, 1e-15, None) == -1
    assert lib.mgs_sketch_reduce(None, 4, 4, 1, 65, None, None, None) == -1


def test_product_does_not_import_the_oracle():
    """The oracle is test infrastructure: nothing under monogs_b200/ (or the drop-in packages) may reference it."""
    for pkg in ("monogs_b200", "diff_gaussian_rasterization", "simple_knn"):
        for dp, _, fs in os.walk(os.path.join(ROOT, pkg)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    txt = open(os.path.join(dp, f)).read()
                    assert "oracle" not in txt.replace("the oracle", "").lower() or f == "scene.py", os.path.join(dp, f)


def test_struct_layouts_match_header(tmp_path):
    """Every struct of the header, compiled as plain C by gcc, has the size and the field offsets of its ctypes mirror in
    monogs_b200/_lib.py (a field added on one side only would shift everything behind it without any error at run time)."""
    import subprocess
    L = _lib()
    pairs = {"mgs_view": L.MgsView, "mgs_loss_args": L.MgsLossArgs, "mgs_adam_group": L.MgsAdamGroup,
             "mgs_track_args": L.MgsTrackArgs, "mgs_track2_args": L.MgsTrack2Args, "mgs_gaussians": L.MgsGaussians,
             "mgs_map_args": L.MgsMapArgs, "mgs_map_view": L.MgsMapView, "mgs_pose_slot": L.MgsPoseSlot, "mgs_peer_exchange": L.MgsPeerExchange, "mgs_densify_cfg": L.MgsDensifyCfg,
             "mgs_gaussian_state": L.MgsGaussianState}
    declared = set(re.findall(r"^}\s*(mgs_\w+);", open(HEADER).read(), flags=re.M))
    assert declared == set(pairs), declared ^ set(pairs)
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {']
    for cname, cls in pairs.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src, exe = tmp_path / "layout.c", tmp_path / "layout"
    src.write_text("\n".join(lines))
    subprocess.run(["gcc", "-std=c11", "-o", str(exe), str(src)], check=True)     # a missing field fails to compile
    got = {}
    for ln in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines():
        c, f, v = ln.split()
        got[(c, f)] = int(v)
    for cname, cls in pairs.items():
        assert got[(cname, "size")] == C.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert got[(cname, fname)] == getattr(cls, fname).offset, (cname, fname)
        # and no field of the header is missing from the mirror: the sizes agree and the last fields line up
