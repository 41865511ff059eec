"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol include/cpe.h declares,
the ctypes mirror agrees with the header, and the product path refuses to run without CUDA (no fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "cpe.h")).read()
    return sorted(set(re.findall(r"^int (cpe_\w+)\(", src, flags=re.M)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from cheetah_pose_estimation_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 9
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_lib.EXPORTS) == names
    assert lib.cpe_version() == 121 == _lib.ABI_VERSION
    hdr = open(os.path.join(ROOT, "include", "cpe.h")).read()
    assert int(re.search(r"#define CPE_VERSION (\d+)", hdr).group(1)) == _lib.ABI_VERSION


def test_struct_mirrors_match_header_field_order():
    from cheetah_pose_estimation_b200 import _lib
    src = open(os.path.join(ROOT, "include", "cpe.h")).read()
    for cname, cls in (("cpe_model", _lib.CpeModel), ("cpe_batch", _lib.CpeBatch), ("cpe_eval_out", _lib.CpeEvalOut), ("cpe_eval_out_f32", _lib.CpeEvalOutF32),
                       ("cpe_assemble_out", _lib.CpeAssembleOut), ("cpe_solve_io", _lib.CpeSolveIO), ("cpe_prior", _lib.CpePrior)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), src, flags=re.S).group(1)
        fields = []
        for line in body.splitlines():
            line = line.split("/*")[0].strip()
            if not line.endswith(";"):
                continue
            decl = re.sub(r"\[[^\]]*\]", "", line[:-1])          # array members: int32_t k2_off[48]
            names = decl.replace("*", " ").split()
            if "," in decl:
                first, rest = decl.split(",", 1)
                fields.append(first.replace("*", " ").split()[-1])
                fields += [r.strip().lstrip("*") for r in rest.split(",")]
            else:
                fields.append(names[-1])
        assert fields == [f[0] for f in cls._fields_], (cname, fields)


def test_argument_errors_are_reported_without_a_gpu():
    from cheetah_pose_estimation_b200 import _lib
    lib = _lib.load()
    assert lib.cpe_eval_f64(None, None, None, None, None) < 0
    assert lib.cpe_band_solve_f64(None, None, None, None, None, None) < 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from cheetah_pose_estimation_b200 import _lib, engine
    with pytest.raises(_lib.CpeError):
        engine.Model("cuda")


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "cheetah_pose_estimation_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert not re.search(r"^\s*(from|import) oracle", open(os.path.join(pkg, fn)).read(), flags=re.M), fn
