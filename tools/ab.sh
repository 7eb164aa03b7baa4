#!/bin/bash
# developer A/B helper (run under gpurun): tools/ab.sh "<lib or empty>" <bench args...>
lib="$1"; shift
BI_B200_LIB="$lib" python bench.py --no-cpu "$@" | python -c "
import json,sys
d=json.loads(sys.stdin.read())
r=d['roofline']; s=d['stages']
print(f\"value {d['value']:.0f} e2e {d['e2e']['value']:.0f} frac {r['frac']:.3f} chol_ms {s['chol_loglik_ms']:.3f} gp_ms {s['gp_predict_ms']:.3f} clk {d['clocks']['sm_mhz']}\")"
