# bench.py against several builds of the library: CG_LIB selects the .so (see conga_b200/abi.py)
for lib in "$@"; do echo "LIB=$lib"; CG_LIB=$lib python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],d['roofline']['frac'],d['roofline']['kernel_ms_per_step'])"; done
