"""Attributes an ncu source-page CSV (SASS rows) to the kernel's phase functions.

usage: python tools/phase_profile.py SRC.csv [kernel-substring]
SASS address -> source line comes from `nvdisasm -g` on the cubin inside
brax_b200/libbrax_b200.so (must be the build that was profiled); a line of
bx_math.cuh is charged to the last bx_step_kernel.cu line seen before it.
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src_csv = sys.argv[1]
kern = sys.argv[2] if len(sys.argv) > 2 else 'ILi0EE'
cu = os.path.join(ROOT, 'brax_b200', 'csrc', 'bx_step_kernel.cu')

# function line ranges of the .cu file
funcs = []
for i, line in enumerate(open(cu), 1):
  m = re.match(r'^(?:template.*\n)?(?:BXD|__device__|extern|__global__|static)[^;]*?\b(\w+)\s*\(', line)
  if m and not line.startswith(' '):
    funcs.append((i, m.group(1)))
  elif re.match(r'^bx_step_kernel\(', line):
    funcs.append((i, 'bx_step_kernel'))
def func_of(ln):
  name = '?'
  for start, n in funcs:
    if start <= ln:
      name = n
  return name

td = tempfile.mkdtemp()
subprocess.check_call(['cuobjdump', '-xelf', 'all', os.path.join(ROOT, 'brax_b200', 'libbrax_b200.so')],
                      cwd=td, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(td) if f.endswith('.cubin')][0]
dis = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(td, cubin)], capture_output=True, text=True).stdout
addr_line, on, cur = {}, False, 0
for l in dis.splitlines():
  if l.startswith('.text.'):
    on = kern in l
  if not on:
    continue
  m = re.search(r'//## File ".*bx_step_kernel\.cu", line (\d+)', l)
  if m:
    cur = int(m.group(1))
  m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/', l)
  if m:
    addr_line[int(m.group(1), 16)] = cur

rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
iA, iE, iS = hdr.index('Address'), hdr.index('Instructions Executed'), hdr.index('# Samples')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
base = None
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot_e = tot_s = 0
for r in data:
  if len(r) < len(hdr):
    continue
  a = int(r[iA], 16) if r[iA].startswith('0x') else int(r[iA])
  if base is None:
    base = a
  f = func_of(addr_line.get(a - base, 0))
  e, s = int(r[iE] or 0), int(r[iS] or 0)
  agg[f][0] += e; agg[f][1] += s
  for i in stall_cols:
    v = int(r[i] or 0)
    if v:
      agg[f][2][hdr[i][6:]] += v
  tot_e += e; tot_s += s
print('%-28s %7s %7s  top stalls' % ('function', 'inst%', 'samp%'))
for f, (e, s, st) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
  print('%-28s %6.1f%% %6.1f%%  %s' % (f, 100 * e / tot_e, 100 * s / max(tot_s, 1),
        ' '.join('%s:%d' % kv for kv in st.most_common(4))))
