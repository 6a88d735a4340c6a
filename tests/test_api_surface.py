"""CPU-side guard: every public entry point resolves to an engine function whose signature accepts
the reference's positional arguments (catches wiring mistakes without a GPU)."""
import inspect

import gadmm_b200 as g
from gadmm_b200 import engine


REF_ARITY = {   # public name -> (engine function, number of positional args of the .m signature)
    "group_ADMM_closedForm": ("gadmm_closed_form", 9),
    "static_group_ADMM_closedForm": ("static_gadmm", 11),
    "dynamic_group_ADMM_closedForm": ("dynamic_gadmm", 12),
    "dynamic_group_ADMM_closedForm_v0": ("dynamic_gadmm_v0", 12),
    "group_ADMM_logistic": ("gadmm_logistic_newton", 10),
    "group_ADMM_logistic_GD": ("gadmm_logistic_gd", 11),
    "standared_ADMM": ("star_admm", 9),
    "GD_DGD_LAG": ("gd_dgd_lag", 10),
    "GD_DGD_LAG_logistic": ("gd_dgd_lag_logistic", 11),
    "dual_averaging": ("dual_averaging", 9),
    "dual_averaging_logisticReg": ("dual_averaging_logistic", 10),
    "opt_sol_closedForm": ("opt_sol_closed_form", 2),
}


def test_public_names_and_engine_targets():
    for name, (target, arity) in REF_ARITY.items():
        assert hasattr(g, name), name
        fn = getattr(engine, target)
        inspect.signature(fn).bind(*range(arity))
        # the public wrapper takes the same positional arguments plus backend=
        sig = inspect.signature(getattr(g, name))
        sig.bind(*range(arity), backend="oracle")
    # the Newton-CG solver behind group_ADMM_logistic(solver="cg") takes the same arguments
    inspect.signature(engine.gadmm_logistic_cg).bind(*range(10), dist=None, return_state=True)
    inspect.signature(engine.gadmm_closed_form).bind(*range(9), clock="simulated")
    for name in ("findPath", "findPath2", "calc_cost", "opt_sol", "opt_sol_logistic", "logReg_GD",
                 "group_ADMM", "standard_ADMM"):
        assert callable(getattr(g, name))


def test_every_family_has_a_torch_distributed_baseline():
    """backend="nccl"/"gloo" covers the chain, star and first-order families (BASELINE.json: the
    NCCL implementation of the same algorithms is the thing to beat)."""
    from gadmm_b200.parallel import dist_backends
    be = dist_backends.get("gloo")
    for name in ("gadmm_closed_form", "static_gadmm", "dynamic_gadmm", "dynamic_gadmm_v0", "star_admm",
                 "gd_dgd_lag", "gd_dgd_lag_logistic", "dual_averaging", "dual_averaging_logistic",
                 "gadmm_logistic", "gadmm_logistic_gd"):
        assert callable(getattr(be, name)), name


def test_engine_classes_expose_their_run_paths():
    from gadmm_b200.engine.chain import ChainEngine
    from gadmm_b200.engine.first_order import FirstOrder
    from gadmm_b200.engine.star import StarEngine
    for cls, names in ((ChainEngine, ("ingest_host", "ingest_gram", "set_path", "iterate", "run",
                                      "run_device_rechain", "persistent_ok", "_run_persistent", "step",
                                      "results", "state", "load_state")),
                       (FirstOrder, ("from_parts", "lambda_max", "hmax_per_worker", "provider",
                                     "run_shared", "run_dgd", "run_dualavg")),
                       (StarEngine, ("iterate", "run", "results"))):
        for n in names:
            assert hasattr(cls, n), (cls.__name__, n)


def test_ctypes_mirrors_match_header_sizes():
    """sizeof of every ctypes mirror equals the C struct size computed from the header layout
    (pointers/doubles/long longs are 8 bytes, ints 4, no padding holes by construction)."""
    import ctypes
    import re
    from gadmm_b200.ops import native as nv
    hdr = open(__file__.replace("tests/test_api_surface.py", "include/gadmm_b200.h")).read()
    for name in ("GadmmChain", "GadmmRechain", "GadmmStar", "GadmmShared", "GadmmOwn", "GadmmLogistic"):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), hdr, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        size = 0
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            n = decl.count(",") + 1
            if "*" in decl or decl.startswith(("double", "long long", "unsigned long long")):
                size += 8 * n
            elif decl.startswith(("int", "unsigned int")):
                size += 4 * n
            else:
                raise AssertionError(f"unparsed member in {name}: {decl!r}")
        size = (size + 7) // 8 * 8
        assert ctypes.sizeof(getattr(nv, name)) == size, (name, ctypes.sizeof(getattr(nv, name)), size)


def test_header_declares_every_exported_entry_point():
    """Every extern "C" function defined in csrc/ is declared in include/gadmm_b200.h, and every
    native call made from Python names a declared function."""
    import glob
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "gadmm_b200.h")).read()
    declared = set(re.findall(r"\b(gadmm_[a-z0-9_]+)\s*\(", hdr))
    defined = set()
    for f in glob.glob(os.path.join(root, "csrc", "*.cu")):
        src = open(f).read()
        if 'extern "C"' not in src:
            continue
        ext = src[src.index('extern "C"'):]
        defined |= set(re.findall(r"^(?:int|size_t|const char\*)\s+(gadmm_[a-z0-9_]+)\s*\(", ext, re.M))
    assert defined and defined <= declared, sorted(defined - declared)
    called = set()
    for f in glob.glob(os.path.join(root, "gadmm_b200", "**", "*.py"), recursive=True):
        called |= set(re.findall(r"(?:lib\(\)|\bL|_lib)\.(gadmm_[a-z0-9_]+)", open(f).read()))
    assert called <= declared, sorted(called - declared)
