"""In-tree build of the native pieces (no JIT cache, so the .so files travel with the repo).

  vllm_metax_b200/libb200_paged_attn.so   C ABI (include/b200_paged_attn.h): CUDA kernels for sm_100a, nvcc only
  vllm_metax_b200/_C.so                   torch-dispatcher shim (csrc/torch_shim.cpp), g++ against torch headers
  oracle/_build/liboracle_ref.so          C restatement of the reference algorithm (test infrastructure)

Run `python -m vllm_metax_b200.build` or call build_all(); everything is incremental on mtimes.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
BUILD = ROOT / "build"
LIB_ABI = PKG / "libb200_paged_attn.so"
LIB_SHIM = PKG / "_C.so"

CU_SOURCES = ["api.cu", "cache_kernels.cu", "attention_generic.cu", "attention_stream.cu", "attention_umma_fp8.cu",
              "merge_attn_states.cu", "cache_movers.cu", "dcp_exchange.cu", "mla_decode.cu", "static_scratch.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA kernels cannot be built")


def _newer(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))


def build_abi(verbose: bool = False) -> Path:
    """nvcc -> libb200_paged_attn.so (sm_100a only)."""
    BUILD.mkdir(exist_ok=True)
    headers = [CSRC / "common.cuh", CSRC / "stream_common.cuh", ROOT / "include" / "b200_paged_attn.h"]
    nvcc = _nvcc()
    jobs = []
    objs = []
    for src in CU_SOURCES:
        obj = BUILD / (src[:-3] + ".o")
        objs.append(obj)
        if _newer(obj, [CSRC / src] + headers):
            extra = ["-DB200_DEBUG"] if os.environ.get("B200_DEBUG") else []
            extra += os.environ.get("B200_EXTRA_NVCC_FLAGS", "").split()  # developer experiments
            jobs.append([nvcc, *NVCC_FLAGS, *extra, "-c", str(CSRC / src), "-o", str(obj)])
    if jobs:
        if verbose:
            print("[build] nvcc x%d" % len(jobs), file=sys.stderr)
        with ThreadPoolExecutor(max_workers=4) as ex:
            list(ex.map(_run, jobs))
    if jobs or _newer(LIB_ABI, objs):
        _run([nvcc, "-shared", "-o", str(LIB_ABI), *map(str, objs), "-cudart", "shared"])
    return LIB_ABI


def build_shim(verbose: bool = False) -> Path:
    """g++ -> _C.so, linked against libb200_paged_attn.so (rpath $ORIGIN) and torch."""
    import torch
    from torch.utils import cpp_extension as ce

    src = CSRC / "torch_shim.cpp"
    if not _newer(LIB_SHIM, [src, ROOT / "include" / "b200_paged_attn.h", LIB_ABI]):
        return LIB_SHIM
    if verbose:
        print("[build] g++ torch_shim.cpp", file=sys.stderr)
    tlib = Path(torch.__file__).parent / "lib"
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI),
           "-DTORCH_EXTENSION_NAME=_C"]
    for inc in ce.include_paths():
        cmd += ["-isystem", inc]
    cmd += ["-isystem", sysconfig.get_paths()["include"], "-isystem", "/usr/local/cuda/include"]
    cmd += [str(src), "-o", str(LIB_SHIM),
            "-L" + str(PKG), "-lb200_paged_attn", "-L" + str(tlib), "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch",
            "-L/usr/local/cuda/lib64", "-lcudart",
            "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + str(tlib), "-Wl,--no-as-needed"]
    _run(cmd)
    return LIB_SHIM


def build_oracle(verbose: bool = False) -> Path | None:
    """gcc -> oracle/_build/liboracle_ref.so (the checker; never loaded by the product path)."""
    src = ROOT / "oracle" / "paged_attn_ref.c"
    if not src.exists():
        return None
    out = ROOT / "oracle" / "_build" / "liboracle_ref.so"
    out.parent.mkdir(exist_ok=True)
    if _newer(out, [src]):
        if verbose:
            print("[build] gcc oracle", file=sys.stderr)
        _run(["gcc", "-O3", "-mavx2", "-mfma", "-fopenmp", "-fPIC", "-shared", "-std=c11", str(src), "-o", str(out), "-lm"])
    return out


def build_all(verbose: bool = False) -> None:
    build_abi(verbose)
    build_shim(verbose)
    build_oracle(verbose)


if __name__ == "__main__":
    build_all(verbose=True)
    print("ok:", LIB_ABI, LIB_SHIM)
