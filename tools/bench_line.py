import sys, json
for l in sys.stdin:
    if l.startswith("{"):
        j = json.loads(l)
        r = j["roofline"]; c = j["clocks"]
        print(f"kernel_ms={r['kernel_ms']:.1f} TF={r['achieved']:.0f} frac={r['frac']:.3f} step_ms={j['ms_per_step']:.1f} "
              f"e2e_ms={j['e2e']['ms_per_step']:.1f} sm_mhz={c['sm_mhz']} pw={c.get('power_w_max')} und={j['config']['undecided_rows_last_step']}")
