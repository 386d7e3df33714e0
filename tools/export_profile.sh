#!/bin/bash
# Export one .ncu-rep (brought back in gpurun_out/) into the text/CSV summaries kept under profiles/.
#   tools/export_profile.sh gpurun_out/prof.ncu-rep profiles/round2_score_kernel
# writes <prefix>_raw.csv (all metrics per profiled launch) and <prefix>_details.txt (the --set full sections),
# and, when the report holds source counters (--import-source on), <prefix>_opcode_mix.txt: executed warp
# instructions per SASS opcode, summed over the first profiled kernel.
set -e
rep=$1; out=$2
ncu -i "$rep" --page raw --csv > "${out}_raw.csv"
ncu -i "$rep" --page details > "${out}_details.txt"
ncu -i "$rep" --page source --csv --print-source sass 2>/dev/null > /tmp/_src.csv || true
python - "$out" <<'PY'
import csv, sys, collections
out = sys.argv[1]
try:
    rows = list(csv.reader(open('/tmp/_src.csv')))
except Exception:
    rows = []
hdr = next((r for r in rows if 'Source' in r and any('Instructions Executed' == c for c in r)), None)
if hdr is None:
    sys.exit(0)
i_src = hdr.index('Source'); i_ex = hdr.index('Instructions Executed')
mix = collections.Counter(); total = 0
start = rows.index(hdr) + 1
for r in rows[start:]:
    if len(r) <= max(i_src, i_ex) or r == hdr:
        break
    try:
        n = int(float(r[i_ex]))
    except ValueError:
        continue
    toks = r[i_src].split()
    op = next((t for t in toks if not t.startswith('@') and not t.startswith('{')), '?').split('.')[0]
    mix[op] += n; total += n
with open(out + '_opcode_mix.txt', 'w') as f:
    f.write(f'# executed warp instructions per SASS opcode, first profiled kernel; total {total}\n')
    for op, n in mix.most_common():
        f.write(f'{op:12s} {n:14d} {100.0 * n / max(total, 1):6.2f} %\n')
PY
ls -la ${out}_*
