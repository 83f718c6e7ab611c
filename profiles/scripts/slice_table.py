"""Tabulate an `ncu --metrics ... --csv` per-launch log of tools/prof_slices.py (developer tool)."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
d = collections.OrderedDict()
for r in rows[hdr + 1:]:
    d.setdefault(r[0], {'grid': r[8]})[r[12]] = float(r[14].replace(',', ''))
print("id grid  us  winst(M) thr/inst  dfma dmul dadd (G)  TF/s")
for k, v in d.items():
    wi = v['smsp__inst_executed.sum']
    ti = v['smsp__thread_inst_executed.sum']
    f = v['sm__sass_thread_inst_executed_op_dfma_pred_on.sum']
    m = v['sm__sass_thread_inst_executed_op_dmul_pred_on.sum']
    a = v['sm__sass_thread_inst_executed_op_dadd_pred_on.sum']
    t = v['gpu__time_duration.sum'] / 1e3
    print(k, v['grid'], f"{t:8.1f} {wi / 1e6:8.1f} {ti / wi:5.1f}  {f / 1e9:6.3f} {m / 1e9:6.3f} {a / 1e9:6.3f}  {(2 * f + m + a) / t / 1e6:6.2f}")
