"""CPU tests of the drop-in boundary: libcoherent_cuda.so loads, exports every symbol that
include/coherent_cuda.h declares, the ctypes mirror matches the header, and the product path
fails loudly -- not silently on some CPU fallback -- when no CUDA device is usable."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "coherent_cuda.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cohcuda_[a-z0-9_]+)\s*\(", src)))


def test_header_compiles_as_c():
    """The boundary is plain C: no C++ or torch types in the signatures."""
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HEADER])


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load_library()
    syms = declared_symbols()
    assert len(syms) >= 20
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    assert sorted(pkg.ABI_SYMBOLS) == syms            # the ctypes mirror lists exactly the header's symbols


def test_struct_layouts_match_header(pkg, tmp_path):
    """sizeof/offsetof of the ABI structs as the C compiler sees them == the ctypes mirror."""
    prog = tmp_path / "lay.c"
    prog.write_text(r'''
#include <stdio.h>
#include <stddef.h>
#include "coherent_cuda.h"
int main(void) {
    printf("%zu %zu %zu %zu %zu %zu\n", sizeof(cohcuda_config), sizeof(cohcuda_chan_state), sizeof(cohcuda_result), sizeof(cohcuda_batch_args),
           sizeof(cohcuda_gather_config), sizeof(cohcuda_gather_geometry));
    printf("%zu %zu %zu %zu %zu\n", offsetof(cohcuda_config, stream), offsetof(cohcuda_batch_args, cf32), offsetof(cohcuda_result, globalseqn),
           offsetof(cohcuda_config, pipe_chunk_blocks), offsetof(cohcuda_gather_config, reserve_sms));
    return 0;
}''')
    exe = tmp_path / "lay"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    a, b = subprocess.check_output([str(exe)], text=True).strip().split("\n")
    sizes = [int(x) for x in a.split()]
    offs = [int(x) for x in b.split()]
    assert sizes == [C.sizeof(pkg.Config), C.sizeof(pkg.ChanState), C.sizeof(pkg.Result), C.sizeof(pkg.BatchArgs),
                     C.sizeof(pkg.GatherConfig), C.sizeof(pkg.GatherGeometry)]
    assert offs == [pkg.Config.stream.offset, pkg.BatchArgs.cf32.offset, pkg.Result.globalseqn.offset,
                    pkg.Config.pipe_chunk_blocks.offset, pkg.GatherConfig.reserve_sms.offset]


def test_no_cpu_fallback(pkg):
    """Without a CUDA device create() must fail with a message; with one it must succeed."""
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        eng = pkg.CoherentCuda(3, 16384)
        eng.close()
        return
    with pytest.raises(pkg.CohCudaError) as e:
        pkg.CoherentCuda(3, 16384)
    assert "no usable CUDA device" in str(e.value) and "no CPU fallback" in str(e.value)


def test_bad_arguments_are_rejected_before_touching_cuda(pkg):
    lib = pkg.load_library()
    h = C.c_void_p()
    for n_signal, B in [(0, 16384), (3, 1000), (3, 32), (70000, 16384)]:
        cfg = pkg.Config(n_signal, B, 0, 0.5, 0, 0, 0, 1, None)
        assert lib.cohcuda_create(C.byref(cfg), C.byref(h)) == -1      # COHCUDA_ERR_ARG
        assert b"blocksize" in lib.cohcuda_last_error(None)
    cfg = pkg.Config(3, 16384, 0, 0.0, 0, 0, 0, 1, None)
    assert lib.cohcuda_create(C.byref(cfg), C.byref(h)) == -1
    assert lib.cohcuda_create(None, C.byref(h)) == -1


def test_product_never_imports_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    pk = os.path.join(ROOT, "coherent-rtlsdr_b200")
    for dirpath, _, files in os.walk(pk):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inc", ".cc", ".cpp", ".hpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f
                assert "cdsp_oracle" not in txt and "orc_" not in txt, f


def test_dropin_library_exports_harness_symbols(oracle):
    """oracle/_ref/libdropin_cuda.so (reference TUs + host/cdsp.cc + host/ccoherent.cc) loads
    without a GPU and resolves libcoherent_cuda.so through its rpath; no compute call is made."""
    import ctypes
    if not oracle.dropin_available():
        pytest.skip("libdropin_cuda.so not built (needs /root/reference at build time)")
    L = ctypes.CDLL(oracle.DROPIN_LIB_PATH)
    for s in ("refh_create", "refh_run_batch", "refh_destroy", "cohcuda_host_prim_handle",
              "cohcuda_host_make_fft_scheme", "cohcuda_host_engine"):
        assert hasattr(L, s), s
