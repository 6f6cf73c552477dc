"""CPU-side checks of the drop-in boundary: libgraphemb.so loads and exports every symbol that
include/graphemb.h declares; no compute call is made (there is no GPU here and no fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from graph_embeddings_b200 import _lib
from graph_embeddings_b200.api import BCA, Configuration, InvalidConfigurationException, Opt

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "graphemb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ge_[a-z0-9_]+)\s*\(", src)) - {"ge_epoch_cb"})


def test_library_exports_every_declared_symbol():
    _lib.build()
    L = C.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/graphemb.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_version_and_default_config_need_no_device():
    L = _lib.lib()
    assert b"sm_100a" in L.ge_version()
    c = _lib.GeConfig()
    L.ge_config_default(C.byref(c))
    # the shipped *.config.yml values (dblp.config.yml:2-3,17-26) and Optimizer.java:26
    assert (c.dim, c.method, c.directed, c.maxiter) == (300, _lib.METHOD["pglove"], 1, 1000)
    assert (c.alpha, c.epsilon, c.tolerance) == (0.1, 1e-3, 1e-4)
    assert abs(c.learning_rate - 0.05) < 1e-9


def test_config_struct_layout_matches_header():
    # field order of ge_config in the header == ctypes mirror
    src = open(os.path.join(ROOT, "include", "graphemb.h")).read()
    body = src[src.index("typedef struct ge_config {"):src.index("} ge_config;")]
    fields = re.findall(r"^\s*(?:int32_t|int64_t|double|float)\s+([a-z_]+);", body, flags=re.M)
    assert fields == [f[0] for f in _lib.GeConfig._fields_]


def test_create_fails_loudly_without_gpu_or_with_bad_config():
    import torch
    L = _lib.lib()
    c = _lib.GeConfig()
    L.ge_config_default(C.byref(c))
    h = C.c_void_p()
    c.dim = 0
    assert L.ge_create(C.byref(c), C.byref(h)) == -1 and b"dimension" in L.ge_last_error(None)
    c.dim, c.alpha = 8, 0.0
    assert L.ge_create(C.byref(c), C.byref(h)) == -1 and b"alpha and epsilon" in L.ge_last_error(None)
    c.alpha = 0.1
    if not torch.cuda.is_available():
        assert L.ge_create(C.byref(c), C.byref(h)) == -7          # GE_ERR_NODEVICE: no CPU fallback
        assert b"no CPU fallback" in L.ge_last_error(None)


def test_configuration_check_mirrors_reference_messages():
    ok = Configuration(graph="g.ttl", method="glove", dim=8, bca=BCA(0.1, 1e-3, True), opt=Opt(), output={"uri": []})
    Configuration.check(ok)
    for mut, msg in [(dict(dim=0), "No dimension specified"), (dict(graph=None), "No input graph specified"),
                     (dict(method=None), "Invalid method"), (dict(bca=BCA(0.0, 1e-3)), "alpha and epsilon are mandatory"),
                     (dict(output=None), "specify at least one group")]:
        bad = Configuration(**{**ok.__dict__, **mut})
        with pytest.raises(InvalidConfigurationException, match=msg):
            Configuration.check(bad)


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import, load or link it, and
    the shared library must not depend on it either."""
    import subprocess
    pkg = os.path.join(ROOT, "graph_embeddings_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), os.path.join(dirpath, f)
                assert "liboracle" not in text, os.path.join(dirpath, f)
    _lib.build()
    needed = subprocess.run(["objdump", "-p", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "NEEDED" in needed and "oracle" not in needed


def test_static_entry_point_allocates_nothing_and_resolves_its_scratch_address():
    """ge_flatten_edges_static (no reference counterpart): its body must not call an allocator, must take the
    device address of the static scratch with a checked cudaGetSymbolAddress, and the header's limits must
    match the array that is compiled in."""
    src = open(os.path.join(ROOT, "graph_embeddings_b200", "csrc", "graph.cu")).read()
    body = src[src.index('extern "C" int ge_flatten_edges_static'):]
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "new ", "malloc("):
        assert banned not in body, banned
    assert "GE_CUDA(cudaGetSymbolAddress((void**)&p, g_flatten_scratch))" in body
    assert "static __device__ __align__(256) unsigned char g_flatten_scratch[ST_BYTES];" in src
    hdr = open(os.path.join(ROOT, "include", "graphemb.h")).read()
    max_e = int(re.search(r"#define GE_STATIC_MAX_EDGES (\d+)", hdr).group(1))
    nbytes = int(re.search(r"#define GE_STATIC_SCRATCH_BYTES (\d+)", hdr).group(1))
    up = lambda b: (b + 255) & ~255
    nb = (max_e + 4095) // 4096
    hist_nb = (256 * nb + 4095) // 4096
    want = 2 * up(max_e * 8) + 3 * up(max_e * 4) + up(256 * nb * 4) + up((max(nb, hist_nb) + 2) * 4 + 16) + 256
    assert nbytes == want and max_e == 1 << 20
    _lib.build()
    elf = subprocess.run(["cuobjdump", "-elf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "g_flatten_scratch" in elf                       # the array is in the device image of the library
