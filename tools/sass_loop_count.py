#!/usr/bin/env python3
"""Static SASS instruction count of the fact walk loop (b7_text<false> inside k_bind_facts) of a libdvb200 build:
sass_loop_count.py lib.so   (needs -lineinfo; counts instructions attributed to the loop's source lines, per line with -v)"""
import re, sys, subprocess, tempfile, os, collections
lib = os.path.abspath(sys.argv[1]); verbose = "-v" in sys.argv
src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "deep_vision_b200", "csrc", "bind7.cuh")).read().split("\n")
def find(p):
    for i, l in enumerate(src):
        if p in l: return i + 1
lo, hi = find("for (int s = 0; s < nsteps; s++) {"), find("if (!CLAIM && B7_RUNG < 3 && ladder")
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "bind.sm_100a.cubin", lib], cwd=d, capture_output=True)
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, "bind.sm_100a.cubin")], capture_output=True, text=True).stdout
fn = cur = None
tot = collections.Counter(); lines = collections.Counter(); ops = collections.Counter()
for line in dis.split("\n"):
    m = re.match(r'\s*\.section\s+\.text\.(\S+?),', line)
    if m: fn = m.group(1); continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s+/\*[0-9a-f]{4,5}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)', line)
    if m and fn:
        tot[fn] += 1
        if 'k_bind_facts' in fn and cur and cur[0] == 'bind7.cuh' and lo <= cur[1] <= hi:
            lines[cur[1]] += 1; ops[m.group(2).split('.')[0]] += 1
print({k.split('10BindParams')[0]: v for k, v in tot.items()})
print("fact walk loop (lines %d-%d): %d instructions" % (lo, hi, sum(lines.values())), ops.most_common(8))
if verbose:
    for l in sorted(lines): print("%4d %3d %s" % (l, lines[l], src[l - 1].strip()[:110]))
