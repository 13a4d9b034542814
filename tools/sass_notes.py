#!/usr/bin/env python3
"""SASS census of the ray-queue traversal kernels (run here, no GPU): instruction mix per kernel from `cuobjdump -sass` of the built
kernels object, and the checks trace_queue.cuh's comments rely on (byte -> float by PRMT, no I2F in the slab test, 128-bit node loads,
no local memory). Usage: python tools/sass_notes.py > profiles/r02_trace_queue_sass_notes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
OBJ = os.path.join(ROOT, "hiprt-path-tracer_b200", "build", "kernels_t.o")
sass = subprocess.run(["cuobjdump", "-sass", OBJ], capture_output=True, text=True, check=True).stdout
usage = subprocess.run(["cuobjdump", "--dump-resource-usage", OBJ], capture_output=True, text=True, check=True).stdout
resources = dict(re.findall(r"Function (\S+):\n\s*(REG:\S+ STACK:\S+ SHARED:\S+)", usage))

functions, name = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        functions[name] = []
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        functions[name].append(m.group(1))

print("SASS census of", os.path.relpath(OBJ, ROOT), "(sm_100a, nvcc", subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout.split("release ")[1].split(",")[0] + ")")
print("columns: instructions | PRMT | I2F* (int -> float conversions) | FFMA+FMUL+FADD | FMNMX* | LDG.*128 | LDG other | LDS/STS | LDL/STL (local memory) | resources")
for fn, ops in functions.items():
    if "k_trace_queue" not in fn:
        continue
    c = collections.Counter(ops)
    total = len(ops)
    def count(pred):
        return sum(v for k, v in c.items() if pred(k))
    short = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip().split("(")[0]
    print(f"{short:60s} {total:5d} | {count(lambda k: k.startswith('PRMT')):4d} | {count(lambda k: k.startswith('I2F')):3d} | "
          f"{count(lambda k: k.split('.')[0] in ('FFMA', 'FMUL', 'FADD')):4d} | {count(lambda k: k.startswith('FMNMX')):4d} | "
          f"{count(lambda k: k.startswith('LDG') and '128' in k):3d} | {count(lambda k: k.startswith('LDG') and '128' not in k):3d} | "
          f"{count(lambda k: k.startswith('LDS') or k.startswith('STS')):3d} | {count(lambda k: k.startswith('LDL') or k.startswith('STL')):3d} | {resources.get(fn, '?')}")
main = [fn for fn in functions if "k_trace_queue_tILb0ELi0" in fn][0]
print("\nfull mnemonic histogram of the default closest-hit kernel (k_trace_queue<false, 0>):")
for k, v in collections.Counter(functions[main]).most_common():
    print(f"  {k:28s} {v}")
