"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol
include/hj_b200.h declares, rejects bad arguments, and fails LOUDLY (HJ_ECUDA, no
fallback) when there is no CUDA device.  No compute calls here."""
import ctypes as C
import json
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_all_exported(hj):
    hdr = open(os.path.join(ROOT, "include", "hj_b200.h")).read()
    declared = set(re.findall(r"\b(hj_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"hj_config", "hj_result", "hj_handle"}
    assert declared == set(hj._capi.ABI_SYMBOLS), declared ^ set(hj._capi.ABI_SYMBOLS)
    for name in hj._capi.ABI_SYMBOLS + hj._capi.HOST_SYMBOLS:
        assert hasattr(hj.lib, name), name


def test_struct_sizes_match_abi(hj):
    cfg = hj.hj_config()
    hj.lib.hj_config_default(C.byref(cfg))
    assert cfg.struct_size == C.sizeof(hj.hj_config)
    # main.cpp:78-85 defaults
    assert (cfg.probe_length, cfg.scale_output, cfg.transaction_size) == (4, 2, 16)
    assert cfg.key_mask == 0xFFFFFFFFFFFFFFFF
    assert hj.lib.hj_abi_version() == 1
    assert C.sizeof(hj.hj_result) % 8 == 0


def test_bad_config_rejected(hj):
    cfg = hj.hj_config()
    hj.lib.hj_config_default(C.byref(cfg))
    h = C.c_void_p()
    cfg.struct_size = 12
    assert hj.lib.hj_create(C.byref(cfg), C.byref(h)) == hj._capi.HJ_EINVAL
    hj.lib.hj_config_default(C.byref(cfg))
    cfg.algo = 17
    assert hj.lib.hj_create(C.byref(cfg), C.byref(h)) == hj._capi.HJ_EINVAL
    hj.lib.hj_config_default(C.byref(cfg))
    cfg.probe_length = 0
    assert hj.lib.hj_create(C.byref(cfg), C.byref(h)) == hj._capi.HJ_EINVAL
    assert hj.lib.hj_create(None, C.byref(h)) == hj._capi.HJ_EINVAL
    assert b"invalid" in hj.lib.hj_strerror(hj._capi.HJ_EINVAL)


def test_no_device_fails_loudly_no_fallback(hj):
    """Without a GPU the product must refuse, not compute on the CPU."""
    if hj.lib.hj_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(hj.HJError) as e:
        hj.HashJoin("atomic")
    assert e.value.code == hj._capi.HJ_ECUDA
    R = hj.generate_data("sorted", 1024)
    with pytest.raises(hj.HJError):
        hj.AtomicHashBuild(R, 1024)


def test_missing_library_is_an_import_error(hj, tmp_path, monkeypatch):
    monkeypatch.setattr(hj._capi, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError):
        hj._capi.load()


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under the package or include/ may name it."""
    bad = []
    for base in ("htm_hashjoin_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", "Makefile")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"liboracle|oracle_lib|orc_[a-z_]+\(|oracle/_ref|oracle\.h", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_cli_flag_behaviour():
    """main.cpp:63-66 (unknown flag -> exit 1), :107-108 (unknown algo -> message, exit 0)."""
    exe = os.path.join(ROOT, "htm_hashjoin_b200", "csrc", "main")
    if not os.path.exists(exe):
        pytest.skip("CLI not built")
    r = subprocess.run([exe, "--bogus", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "Unknown option" in r.stdout
    r = subprocess.run([exe, "--algo", "foo"], capture_output=True, text=True)
    assert r.returncode == 0 and "Unknown Algo: foo" in r.stdout
    r = subprocess.run([exe, "--algo", "atomic", "--rSize", "1K", "--dataDistr", "zipf"], capture_output=True,
                       text=True)
    assert r.returncode == 1 and "Unknown distribution" in r.stdout


def test_header_enums_match_the_python_binding(hj):
    """HJ_ALGO_* / HJ_PATH_* / HJ_E* in include/hj_b200.h are what _capi.py binds (incl. the sort-merge comparator)."""
    hdr = open(os.path.join(ROOT, "include", "hj_b200.h")).read()
    vals = {k: int(v) for k, v in re.findall(r"\b(HJ_(?:ALGO|PATH|E)[A-Z_]*|HJ_OK)\s*=\s*(-?\d+)", hdr)}
    assert vals["HJ_ALGO_SORTMERGE"] == 4 and vals["HJ_PATH_SORTMERGE"] == 3
    for name, v in vals.items():
        if hasattr(hj._capi, name):
            assert getattr(hj._capi, name) == v, name
    assert hj.ALGOS["sortmerge"] == vals["HJ_ALGO_SORTMERGE"]
    cfg = hj.hj_config()
    hj.lib.hj_config_default(C.byref(cfg))
    cfg.algo = vals["HJ_ALGO_SORTMERGE"] + 1  # one past the last algorithm
    h = C.c_void_p()
    assert hj.lib.hj_create(C.byref(cfg), C.byref(h)) == hj._capi.HJ_EINVAL


def test_cli_sortmerge_is_a_known_algo():
    """SortMerge.cpp is its own binary in the reference; here it is --algo sortmerge.  It needs a probe side, and
    without a GPU it reports the error in the JSON line (exit 2) instead of computing on the CPU."""
    exe = os.path.join(ROOT, "htm_hashjoin_b200", "csrc", "main")
    if not os.path.exists(exe):
        pytest.skip("CLI not built")
    r = subprocess.run([exe, "--algo", "sortmerge", "--rSize", "1K", "--probe", "0"], capture_output=True, text=True)
    assert r.returncode == 1 and "needs a probe side" in r.stdout
    r = subprocess.run([exe, "--algo", "sortmerge", "--rSize", "1K"], capture_output=True, text=True)
    assert "Unknown Algo" not in r.stdout
    lib = C.CDLL(os.path.join(ROOT, "htm_hashjoin_b200", "csrc", "libhj_b200.so"))
    if lib.hj_device_count() <= 0:
        assert r.returncode == 2 and '"error"' in r.stdout and "no CPU fallback" in r.stdout


def test_reference_main_cpp_compiles_against_ref_entry(tmp_path):
    """INTEGRATION.md section A, executed: the reference's own main.cpp with the three operator includes
    (main.cpp:10-14) swapped for ref_entry.hpp compiles UNCHANGED otherwise (both ENABLE_PROBE settings)
    and links against libhj_b200.so; without a GPU the run prints the JSON error line (no CPU fallback)."""
    import re
    import subprocess
    ref = "/root/reference"
    if not os.path.exists(os.path.join(ref, "main.cpp")):
        pytest.skip("reference checkout not present (GPU box)")
    src = open(os.path.join(ref, "main.cpp")).read()
    swapped, n = re.subn(r'#include "(AtomicHashBuild|HTMHashBuild|NoCCHashBuild)\.hpp"', "", src)
    assert n == 3
    swapped = swapped.replace('#include "config.h"', '#include "config.h"\n#include "ref_entry.hpp"', 1)
    csrc = os.path.join(ROOT, "htm_hashjoin_b200", "csrc")
    for probe in (0, 1):
        cfg_dir = tmp_path / f"cfg{probe}"
        cfg_dir.mkdir()
        cfg = open(os.path.join(ref, "config.h")).read()
        (cfg_dir / "config.h").write_text(re.sub(r"#define ENABLE_PROBE \d", f"#define ENABLE_PROBE {probe}", cfg))
        (cfg_dir / "main_swapped.cpp").write_text(swapped)
        exe = cfg_dir / "main"
        cmd = ["g++", "-std=c++14", "-O1", "-w", "-include", "cstdint", f"-I{cfg_dir}", f"-I{ROOT}/oracle/tbb_shim",
               f"-I{csrc}", f"-I{ref}", str(cfg_dir / "main_swapped.cpp"), "-o", str(exe), f"-L{csrc}", "-lhj_b200",
               f"-Wl,-rpath,{csrc}", "-pthread"]
        out = subprocess.run(cmd, capture_output=True, text=True)
        assert out.returncode == 0, out.stderr[-2000:]
        run = subprocess.run([str(exe), "--algo", "atomic", "--rSize", "1024", "--dataDistr", "sorted"],
                             capture_output=True, text=True, timeout=120)
        line = json.loads(run.stdout.strip().splitlines()[-1])
        assert line["algo"] == "atomic" and line["rSize"] == 1024
        try:
            import torch
            has_gpu = torch.cuda.is_available()
        except Exception:
            has_gpu = False
        if has_gpu:
            assert line["inputSum"] == line["outputSum"] == 1024 * 1025 // 2
        else:
            assert "error" in line and "hj_create" in line["error"]
