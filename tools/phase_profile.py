"""Attributes an ncu source-page CSV (SASS rows) to the kernel's phase functions.

usage: python tools/phase_profile.py SRC.csv KERNEL_REGEX [--lines N] [--ops]
  KERNEL_REGEX selects the .text section, e.g. 'ILi0ELi\\d+ELi320ELi2E' ; with
  several matches the candidates are listed.
SASS address -> source line comes from `nvdisasm -g` on the cubins inside
brax_b200/libbrax_b200.so (must be the build that was profiled); a line of
bx_math.cuh is charged to the last bx_step_kernel.cuh line seen before it.
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
kern = re.compile(sys.argv[2])
nlines = int(sys.argv[sys.argv.index('--lines') + 1]) if '--lines' in sys.argv else 0
cuh = os.path.join(ROOT, 'brax_b200', 'csrc', 'bx_step_kernel.cuh')
src_lines = open(cuh).read().splitlines()

funcs = []
for i, line in enumerate(src_lines, 1):
  m = re.match(r'^(?:BXS|BXD|__device__|static)\s[^;]*?\b(\w+)\s*\($', line) or \
      re.match(r'^(?:BXS|BXD|__device__|static)\s[^;]*?\b(\w+)\s*\(', line)
  if m:
    funcs.append((i, m.group(1)))


math_h = os.path.join(ROOT, 'brax_b200', 'csrc', 'bx_math.cuh')
mfuncs = []
for i, line in enumerate(open(math_h).read().splitlines(), 1):
  m = re.match(r'^(?:BXD|__device__|static)\s[^;]*?\b(\w+)\s*\(', line)
  if m:
    mfuncs.append((i, m.group(1)))


def mfunc_of(ln):
  name = '?'
  for start, n in mfuncs:
    if start <= ln:
      name = n
  return 'math:' + name


def func_of(ln):
  name = '?'
  for start, n in funcs:
    if start <= ln:
      name = n
  return name


# the kernel objects of the last build (brax_b200/build.py keeps them in the temp dir);
# every launch-bound TU has its own cubin, extracted into its own directory
objdir = os.path.join(tempfile.gettempdir(), 'brax_b200_obj_libbrax_b200.so')
cubins = []
for o in sorted(f for f in os.listdir(objdir) if f.startswith('bx_kernels_') and f.endswith('.o')):
  td = tempfile.mkdtemp()
  subprocess.check_call(['cuobjdump', '-xelf', 'all', os.path.join(objdir, o)], cwd=td, stdout=subprocess.DEVNULL)
  cubins += [os.path.join(td, f) for f in os.listdir(td) if f.endswith('.cubin')]
addr_line, addr_op, addr_math, found = {}, {}, {}, []
curm = None
for cubin in cubins:
  dis = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout
  on, cur = False, 0
  for l in dis.splitlines():
    if l.startswith('.text.'):
      on = bool(kern.search(l))
      if on:
        found.append(l.strip())
    if not on:
      continue
    m = re.search(r'//## File ".*bx_step_kernel\.cuh", line (\d+)', l)
    if m:
      cur = int(m.group(1))
      curm = None
    m = re.search(r'//## File ".*bx_math\.cuh", line (\d+)', l)
    if m:
      curm = mfunc_of(int(m.group(1)))
    m = re.match(r'\s+/\*([0-9a-f]{4,6})\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', l)
    if m:
      addr_line[int(m.group(1), 16)] = cur
      addr_op[int(m.group(1), 16)] = m.group(2)
      addr_math[int(m.group(1), 16)] = curm
if len(found) != 1:
  print('kernel regex matched %d sections:' % len(found))
  for f in found[:20]:
    print(' ', f)
  sys.exit(1)
print('kernel', found[0])

rows = list(csv.reader(open(src_csv)))
hdr, data = rows[1], rows[2:]
iA, iE, iS = hdr.index('Address'), hdr.index('Instructions Executed'), hdr.index('# Samples')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
base = None
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
by_line = collections.defaultdict(lambda: [0, 0])
ops = collections.Counter()
tot_e = tot_s = 0
for r in data:
  if len(r) < len(hdr):
    continue
  a = int(r[iA], 16) if r[iA].startswith('0x') else int(r[iA])
  if base is None:
    base = a
  ln = addr_line.get(a - base, 0)
  f = func_of(ln)
  e, s = int(r[iE] or 0), int(r[iS] or 0)
  agg[f][0] += e; agg[f][1] += s
  by_line[ln][0] += e; by_line[ln][1] += s
  ops[addr_op.get(a - base, '?').split('.')[0]] += e
  for i in stall_cols:
    v = int(r[i] or 0)
    if v:
      agg[f][2][hdr[i][6:]] += v
  tot_e += e; tot_s += s
print('%-28s %7s %7s  top stalls' % ('function', 'inst%', 'samp%'))
for f, (e, s, st) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
  print('%-28s %6.1f%% %6.1f%%  %s' % (f, 100 * e / tot_e, 100 * s / max(tot_s, 1),
        ' '.join('%s:%d' % kv for kv in st.most_common(4))))
if '--static' in sys.argv:
  st = collections.Counter(addr_math.get(a) or func_of(l) for a, l in addr_line.items())
  print('\nstatic SASS instructions by function (total %d = %d KB):' % (len(addr_line), len(addr_line) * 16 // 1024))
  for f, n in st.most_common(40):
    print('  %-28s %5d  %5.1f KB' % (f, n, n * 16 / 1024))
if '--ops' in sys.argv:
  print('\nopcode mix (warp instructions executed):')
  for op, n in ops.most_common(24):
    print('  %-10s %5.1f%%' % (op, 100 * n / tot_e))
if nlines:
  print('\ntop source lines by samples:')
  for ln, (e, s) in sorted(by_line.items(), key=lambda kv: -kv[1][1])[:nlines]:
    text = src_lines[ln - 1].strip() if 0 < ln <= len(src_lines) else ''
    print('  %5d %-22s inst %4.1f%% samp %4.1f%%  %s' % (ln, func_of(ln), 100 * e / tot_e, 100 * s / max(tot_s, 1), text[:70]))
