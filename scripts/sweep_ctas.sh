for n in ${CTAS_LIST:-5 4 3 2}; do echo "CTAS=$n"; CG_PASS_CTAS=$n python scripts/pass_bench.py 2>&1 | grep "^{"; CG_PASS_CTAS=$n python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],d['roofline']['frac'],d['roofline']['kernel_ms_per_step'])"; done
