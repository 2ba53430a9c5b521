for n in ${LIST:-0 64 32 16 8}; do echo "GENO_BLOCKS=$n"; CG_GENO_BLOCKS=$n python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],d['roofline']['frac'],d['roofline']['kernel_ms_per_step'])"; done
