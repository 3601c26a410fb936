"""Per source function: warp instructions executed per bin and stall samples, from an `ncu --set full --import-source on` SASS
source page (CSV, possibly .gz) and the line info of the very build that ran (nvdisasm --print-line-info of the cubin).

    cuobjdump -xelf all build/obj/avr_lib.cu.o && nvdisasm --print-line-info avr_lib.sm_100a.cubin > lines.txt
    python tools/ncu_by_function.py gpurun_out/r3g_source_sass.csv.gz lines.txt CompressELb1 18418681

(function names are the innermost inlined function's; several classes share `put` / `run`)"""
import bisect, collections, csv, gzip, io, os, re, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src_csv, lines_txt, kernel_tag, bins = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
sec = cur = None
amap = {}
for ln in open(lines_txt):
    m = re.match(r'^\s*\.section\s+\.text\.(\S+?),', ln)
    if m:
        sec, cur = m.group(1), None
        continue
    if sec is None or kernel_tag not in sec:
        continue
    m = re.match(r'^\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'^\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', ln)
    if m:
        amap[int(m.group(1), 16)] = cur
funcs = {}
for fn in ('avr_core.cuh', 'avr_walk.cuh', 'avr_job.cuh', 'avr_lib.cu'):
    starts = []
    for i, l in enumerate(open(os.path.join(ROOT, 'avrecode_b200', 'csrc', fn)), 1):
        m = re.match(r'\s*(?:template <[^>]*>\s*)?(?:AVR_DEV_NOINLINE|AVR_DEV|__global__|inline|static)\b[^;=]*?\b(\w+)\s*\(', l)
        if m and not l.strip().startswith('//'):
            starts.append((i, m.group(1)))
    funcs[fn] = starts


def func_of(key):
    if key is None:
        return 'unknown'
    fn, line = key
    st = funcs.get(fn)
    if not st:
        return fn
    idx = bisect.bisect_right([s[0] for s in st], line) - 1
    return fn.split('.')[0][4:] + ':' + (st[idx][1] if idx >= 0 else '?')


opener = gzip.open if src_csv.endswith('.gz') else open
rows = list(csv.reader(io.StringIO(opener(src_csv, 'rt').read())))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'Address'] + [len(rows)]
for k in range(len(hdr) - 1):
    h = rows[hdr[k]]
    col = {n: i for i, n in enumerate(h)}
    body = [r for r in rows[hdr[k] + 1:hdr[k + 1]] if len(r) == len(h)]
    if len(body) != len(amap):
        continue   # another kernel of the report
    base = int(body[0][col['Address']], 16)

    def f(r, n):
        try:
            return float(r[col[n]])
        except ValueError:
            return 0.0
    agg = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0, 0])
    for r in body:
        a = agg[func_of(amap.get(int(r[col['Address']], 16) - base))]
        a[0] += f(r, 'Instructions Executed'); a[1] += f(r, '# Samples'); a[2] += f(r, 'stall_no_inst'); a[3] += f(r, 'stall_branch_resolving'); a[4] += 1
    tot = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
    print("kernel *%s*: %d instructions, %.1f warp instructions executed per bin (all warps, waits included)" % (kernel_tag, len(body), tot / bins))
    for name, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
        print("%-30s sass %5d  exec/bin %6.2f  samples %5.2f%%  of them: instruction fetch %4.1f%% branch %4.1f%%" %
              (name, v[4], v[0] / bins, 100 * v[1] / ts, 100 * v[2] / max(v[1], 1), 100 * v[3] / max(v[1], 1)))
