"""profiles/rNN_sweep_config5.jsonl (scripts/sweep_config5.py) -> markdown rows:  python scripts/sweep_table.py FILE..."""
import json
import sys

print("| P | D | T | ms | latents/s | TFLOP/s | frac of peak | rel-L2 | index mismatches | failed certificate / full fp32 pass |")
print("|---|---|---|---|---|---|---|---|---|---|")
for f in sys.argv[1:]:
    for line in open(f):
        d = json.loads(line)
        print(f"| {d['patterns']} | {d['dim']} | {d['update_steps']} | {d['ms']:.3f} | {d['latents_per_s']:.3g} | "
              f"{d['algorithmic_tflops']:.0f} | {d['frac_of_measured_peak']:.3f} | {d['rel_l2']:.2e} | "
              f"{d['argmax_mismatch']} / {d['argmax_decided_rows']} | {d.get('rows_failed_certificate', '')} / {d.get('rows_rescored_fp32', '')} |")
