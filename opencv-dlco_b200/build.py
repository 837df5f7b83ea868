"""Builds libdlco_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdlco_b200.so")
SOURCES = ["api.cu", "desc.cu", "pool_simt.cu", "pool_tc.cu", "fulldist.cu", "hinge.cu",
           "pr_learn.cu", "pj_learn.cu", "stats.cu", "gemm_tc.cu", "sqnorm_tc.cu", "pj_sign.cu", "gen_bank.cu", "comm.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(HERE, "..", "include", "dlco_b200.h"))
    objs, procs = [], []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        if not os.path.exists(s):
            continue
        o = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + headers):
            cmd = [nvcc] + NVCC_FLAGS + ["-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd))
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src, out.decode()))
        if verbose and out:
            print(out.decode())
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s" % r.stdout.decode())
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))


TOOLS = ["comp-fulldists", "comp-uprjdists", "pr-learn", "pj-learn", "pj-sweep", "pr-stats", "pj-stats", "h5tool",
         "gen-poolregion", "export-opencv", "conv-impatches"]


def build_tools(force=False, verbose=False):
    """The drop-in command-line tools (tools/*.cpp, C++ host code over the C ABI) -> bin/."""
    root = os.path.abspath(os.path.join(HERE, ".."))
    src_dir, bin_dir = os.path.join(root, "tools"), os.path.join(root, "bin")
    os.makedirs(bin_dir, exist_ok=True)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    headers = [os.path.join(src_dir, f) for f in os.listdir(src_dir) if f.endswith(".hpp")]
    headers.append(os.path.join(root, "include", "dlco_b200.h"))
    procs = []
    for t in TOOLS:
        src = os.path.join(src_dir, t + ".cpp")
        exe = os.path.join(bin_dir, t)
        if not os.path.exists(src):
            continue
        if not (force or _stale(exe, [src] + headers)):
            continue
        cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-I", os.path.join(cuda, "include"), src, "-o", exe]
        if t not in ("h5tool", "export-opencv", "conv-impatches"):
            cmd += ["-L", HERE, "-ldlco_b200", "-L", os.path.join(cuda, "lib64"), "-lcudart",
                    "-Wl,-rpath,$ORIGIN/../opencv-dlco_b200", "-Wl,-rpath," + os.path.join(cuda, "lib64")]
        cmd += ["-lz", "-lpthread"]
        if verbose:
            print(" ".join(cmd))
        procs.append((t, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for t, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("g++ failed on tools/%s.cpp:\n%s" % (t, out.decode()))
        if verbose and out:
            print(out.decode())
    return bin_dir
