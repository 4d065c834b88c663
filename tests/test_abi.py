"""CPU tests of the drop-in boundary: libkmi.so loads, exports exactly what include/kmi.h declares,
and refuses to compute without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re
import subprocess

import pytest

import kmerind_b200 as kb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(kb.LIB_PATH):
        from kmerind_b200 import build
        build.build()
    return kb.load_library()


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "kmi.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(kmi_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(lib):
    decl = declared_symbols()
    assert len(decl) >= 30
    for name in decl:
        assert hasattr(lib, name), "libkmi.so does not export %s" % name
    assert sorted(kb.EXPORTED_SYMBOLS) == decl
    out = subprocess.run(["nm", "-D", "--defined-only", kb.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r" T (kmi_[a-z0-9_]+)", out)))
    assert exported == decl, "exported C symbols differ from include/kmi.h"


def test_header_compiles_as_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "kmi.h"\nint main(void){kmi_config c; c.k=31; return (int)sizeof(c) - 40;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o",
                           str(tmp_path / "t")])
    assert subprocess.run([str(tmp_path / "t")]).returncode == 0
    assert ctypes.sizeof(kb.KmiConfig) == 40


def test_cpp_adapter_compiles(tmp_path):
    hdr = os.path.join(ROOT, "include", "kmi_index.hpp")
    if not os.path.exists(hdr):
        pytest.skip("adapter header not present")
    src = tmp_path / "t.cpp"
    src.write_text('#include "kmi_index.hpp"\n#include <vector>\nint main(){ if (false) { kmi::CountIndex<31> idx; '
                   'std::vector<kmi::Kmer<31>> v; idx.insert(v); auto r = idx.find(v); auto c = idx.count(v); idx.erase(v);'
                   ' (void)r; (void)c; (void)idx.size(); (void)idx.local_size(); } return 0; }\n')
    subprocess.check_call(["g++", "-std=c++11", "-Wall", "-I", os.path.join(ROOT, "include"), str(src), kb.LIB_PATH, "-o",
                           str(tmp_path / "t"), "-Wl,-rpath," + os.path.dirname(kb.LIB_PATH)])


REFERENCE = "/root/reference"


@pytest.mark.parametrize("flavour", ["kmi", "bliss"])
def test_adapter_check_compiles_with_both_key_types(flavour, tmp_path):
    """tests/cpp/adapter_check.cpp (every member Index<MapType,KmerParser> forwards to, kmer_index.hpp:115-194) instantiates
    kmi::b200_map with kmi::Kmer<K> and — where the reference tree is present — with the reference's own
    bliss::common::Kmer<K, DNA, uint64_t> (private data, no canonical() member) and a comm type that only has rank()/size(),
    as mxx::comm does. Compile + link only: running it needs a GPU (tests/test_gpu_parity.py)."""
    cmd = ["g++", "-std=c++11", "-O0", "-w", "-I", os.path.join(ROOT, "include")]
    if flavour == "bliss":
        if not os.path.isdir(os.path.join(REFERENCE, "src")):
            pytest.skip("reference tree not present (GPU box): the prebuilt oracle/_ref/adapter_check_bliss is run instead")
        cmd += ["-DKMI_CHECK_BLISS", "-DNDEBUG", "-I", os.path.join(ROOT, "oracle", "refcfg"), "-I", os.path.join(REFERENCE, "src"),
                "-I", os.path.join(REFERENCE, "ext")]
    cmd += [os.path.join(ROOT, "tests", "cpp", "adapter_check.cpp"), kb.LIB_PATH, "-o", str(tmp_path / "t"),
            "-Wl,-rpath," + os.path.dirname(kb.LIB_PATH)]
    subprocess.check_call(cmd)


def test_argument_validation_and_no_cpu_fallback(lib):
    import torch
    h = ctypes.c_void_p()
    for bad_k in (0, 1, 32, 64, 100):
        cfg = kb.KmiConfig(bad_k, 0, 1, 0, 0, 0, 0, 1, 0)
        assert lib.kmi_create(ctypes.byref(cfg), ctypes.byref(h)) == kb.ERR_ARG
    # K outside the built-in set needs its kernel module (kmerind_b200/kmods/libkmi_k<K>.so): 34 is never built by default
    cfg = kb.KmiConfig(34, 0, 1, 0, 0, 0, 0, 1, 0)
    assert lib.kmi_create(ctypes.byref(cfg), ctypes.byref(h)) == kb.ERR_UNSUPPORTED
    assert b"kmods/libkmi_k34.so" in lib.kmi_last_error()
    cfg = kb.KmiConfig(31, 0, 1, 7, 0, 0, 0, 1, 0)
    assert lib.kmi_create(ctypes.byref(cfg), ctypes.byref(h)) == kb.ERR_ARG
    cfg = kb.KmiConfig(31, 0, 1, 0, 0, 0, 2, 2, 0)
    assert lib.kmi_create(ctypes.byref(cfg), ctypes.byref(h)) == kb.ERR_ARG
    if not torch.cuda.is_available():
        cfg = kb.KmiConfig(31, 0, 1, 0, 0, 0, 0, 1, 0)
        assert lib.kmi_create(ctypes.byref(cfg), ctypes.byref(h)) == kb.ERR_CUDA
        assert b"no CPU path" in lib.kmi_last_error()
        with pytest.raises(kb.KmiError):
            kb.CountIndex(31)
    assert lib.kmi_synth_record_bytes(150) == 317
