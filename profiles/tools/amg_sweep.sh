mkdir -p gpurun_out/s3m
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/s3m/bench_n1.json 2> gpurun_out/s3m/bench_n1.err
python -c "
import json;d=json.loads(open('gpurun_out/s3m/bench_n1.json').read().strip().splitlines()[-1]);print('default',d['value'],d['ms_per_step'],d['phases_ms'],'e2e',d['e2e']['value'],d['e2e']['breakdown_s'])"
for cfg in "MV_AMG_DEGC=1" "MV_AMG_THETA=0.05" "MV_AMG_THETA=0.1" "MV_AMG_DEGC=1 MV_AMG_THETA=0.05" "MV_AMG_DEG0=2"; do
  env $cfg timeout 200 python bench.py --steps 10 --warmup 5 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/s3m/b.json 2>/dev/null
  python -c "
import json,sys;d=json.loads(open('gpurun_out/s3m/b.json').read().strip().splitlines()[-1]);print(sys.argv[1],d['value'],d['ms_per_step'],d['phases_ms']['poisson_pseudotimestep(A7)'],d['config']['inner_iterations_per_step(l2,poisson,hessian)'],d['config']['poisson_solver'])" "$cfg"
done
