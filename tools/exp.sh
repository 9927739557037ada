for cfg in "7 0" "5 0" "7 1"; do set -- $cfg
  export SOMAR_B200_TMA_PROD=$1; if [ "$2" = "1" ]; then export SOMAR_B200_TMA_NOLOAD=1; else unset SOMAR_B200_TMA_NOLOAD; fi
  timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_exp.json 2> gpurun_out/bench_exp.err; tail -2 gpurun_out/bench_exp.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_exp.json'));print('prod/noload $cfg', d['value'], d['roofline']['avg_launch_ms'])"
done
