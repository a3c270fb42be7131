"""Stall samples and executed instructions of one kernel by CUDA source line (reads an ncu report captured with
--set full --import-source on).  usage: python profiles/by_source_line.py <report.ncu-rep> [n_lines]"""
import csv, subprocess, sys

rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
cur, hdr, agg = None, None, []
for r in csv.reader(out.splitlines()):
    if r and r[0] == 'File Path':
        cur = r[1].split('/')[-1]
    elif r and r[0] == 'Line No':
        hdr = r
    elif hdr is not None and len(r) >= len(hdr) and r[2] == '-':      # a CUDA line (SASS rows carry an address)
        agg.append((cur, int(r[0]), r[1], r))
ix = {n: i for i, n in enumerate(hdr)}
S, I = ix['# Samples'], ix['Instructions Executed']
tot_s, tot_i = sum(int(a[3][S]) for a in agg), sum(int(a[3][I]) for a in agg)
stalls = [n for n in hdr if n.startswith('stall_') and 'Not Issued' not in n]
print(f'# {rep}: {tot_s} stall samples, {tot_i} warp instructions (inlined callees are attributed to their own lines)')
by_reason = sorted(((sum(int(a[3][ix[n]]) for a in agg), n[6:]) for n in stalls), reverse=True)
print('# samples by reason: ' + ', '.join(f'{n} {100 * v / tot_s:.1f}%' for v, n in by_reason[:9]))
agg.sort(key=lambda a: -int(a[3][S]))
for f, l, src, r in agg[:top]:
    st = sorted(((int(r[ix[n]]), n[6:]) for n in stalls), reverse=True)[:3]
    print(f'{f[:12]:12s}:{l:5d} {100 * int(r[S]) / tot_s:5.1f}% smp {100 * int(r[I]) / tot_i:5.1f}% ins  '
          f'{" ".join(f"{n}={v}" for v, n in st):42s} {src.strip()[:88]}')
