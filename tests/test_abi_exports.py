"""The C-ABI library loads and exports every entry point include/oe_abi.h declares (no compute calls: runs without a GPU)."""
import ctypes
import os
import re

from dcos_commons_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DECL = re.compile(r"^(?:int32_t|int64_t|void|const char\*|void\*)\s+(oe_\w+)\s*\(", re.M)


def declared_functions():
    text = open(os.path.join(ROOT, "include", "oe_abi.h")).read()
    return sorted(set(DECL.findall(text)))


def test_header_declares_the_expected_surface():
    names = declared_functions()
    for must in ("oe_create", "oe_destroy", "oe_last_error", "oe_load_programs", "oe_load_tasks", "oe_load_offers", "oe_load_requirements",
                 "oe_feasibility", "oe_class_feasibility", "oe_evaluate_cycle", "oe_apply_launches", "oe_export_tasks", "oe_device_buffer",
                 "oe_stream", "oe_matrix_stride_words", "oe_abi_version"):
        assert must in names


def test_engine_library_exports_every_declared_symbol(engine_lib):
    lib = ctypes.CDLL(abi.ENGINE_LIB)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, f"liboe_b200.so lacks {missing}"
    assert engine_lib.fn("abi_version")() == abi.OE_ABI_VERSION


def test_struct_sizes_match_the_header():
    assert abi.STAGE_OP_DTYPE.itemsize == 32 and abi.RULE_NODE_DTYPE.itemsize == 16 and abi.DECISION_DTYPE.itemsize == 32
    assert ctypes.sizeof(abi.OeConfig) == 24 and ctypes.sizeof(abi.OeDictSizes) == 48 and ctypes.sizeof(abi.OeStats) == 48


def test_without_a_gpu_the_engine_refuses_to_start_instead_of_falling_back():
    import torch
    if torch.cuda.is_available():
        return
    try:
        abi.Engine(device=0)
    except abi.EngineError as e:
        assert e.code == -4 and "no CPU path" in str(e)
    else:
        raise AssertionError("an engine came up without a CUDA device")


def test_oracle_exports_the_same_entry_points_under_its_own_prefix(oracle_lib):
    for n in ("create", "destroy", "load_programs", "load_tasks", "load_offers", "load_requirements", "feasibility", "evaluate_cycle",
              "apply_launches", "export_tasks", "matrix_stride_words"):
        assert oracle_lib.has(n)


def test_header_is_plain_c_and_a_c_program_links_against_the_library(tmp_path, engine_lib):
    """include/oe_abi.h must be usable from C (what a JNI shim is written in); the program is built, and run only as far as
    oe_create, which refuses cleanly on a box without a GPU."""
    import subprocess
    exe = tmp_path / "c_abi_link"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-o", str(exe), os.path.join(ROOT, "tests", "c_abi_link.c"),
                           abi.ENGINE_LIB, "-Wl,-rpath," + os.path.dirname(abi.ENGINE_LIB)])
    import torch
    rc = subprocess.run([str(exe)]).returncode
    assert rc == 0 if torch.cuda.is_available() else rc in (0, 1), rc
