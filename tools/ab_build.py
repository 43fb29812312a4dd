"""Builds a variant of libmjb200.so with extra -D flags into gpurun_out-independent path csrc/variants/<name>.so
(git-ignored, travels to the GPU box).  usage: ab_build.py name -DMJB_SOL_MIN_CTAS=4 ...
Select at run time with MJB200_LIB=mujoco_warp_b200/csrc/variants/<name>.so"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "mujoco_warp_b200", "csrc")
out = os.path.join(CSRC, "variants")
os.makedirs(out, exist_ok=True)
name, flags = sys.argv[1], sys.argv[2:]
dst = os.path.join(out, name + ".so")
cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xptxas", "-v", "-shared", "-Xcompiler", "-fPIC",
       "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-o", dst, os.path.join(CSRC, "mjb_api.cu")] + flags
r = subprocess.run(cmd, capture_output=True, text=True)
if r.returncode:
  sys.exit(r.stdout + r.stderr)
import re
for m in re.finditer(r"Compiling entry function '(\S+)'.*?\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores.*?\n.*?Used (\d+) registers", r.stderr):
  if "group_kernel" in m.group(1) and "Lb1" in m.group(1):
    print(m.group(1), "regs", m.group(4), "spill", m.group(3))
print("built", dst)
