"""Builds a variant of libmjb200.so with extra -D flags into mujoco_warp_b200/csrc/variants/<name>.so (git-ignored, travels to
the GPU box); every translation unit is recompiled, in parallel.  usage: ab_build.py name -DMJB_SOL_MIN_CTAS=4 ...
Select at run time with MJB200_LIB=mujoco_warp_b200/csrc/variants/<name>.so"""
import os, subprocess, sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "mujoco_warp_b200", "csrc")
out = os.path.join(CSRC, "variants")
name, flags = sys.argv[1], sys.argv[2:]
objdir = os.path.join(out, "_obj_" + name)
os.makedirs(objdir, exist_ok=True)
dst = os.path.join(out, name + ".so")
srcs = [f for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]
common = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
          "-I", os.path.join(ROOT, "include"), "-I", CSRC] + flags


def one(f):
  o = os.path.join(objdir, f[:-3] + ".o")
  r = subprocess.run(common + ["-c", os.path.join(CSRC, f), "-o", o], capture_output=True, text=True)
  if r.returncode:
    sys.exit(r.stdout + r.stderr)
  return o


with ThreadPoolExecutor(max_workers=os.cpu_count()) as pool:
  objs = list(pool.map(one, srcs))
r = subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", dst] + objs, capture_output=True, text=True)
if r.returncode:
  sys.exit(r.stdout + r.stderr)
print("built", dst)
