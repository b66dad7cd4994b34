#!/bin/bash
# One gpurun call of the round: tools/gpu_call.sh <tag> <steps...>   (results under gpurun_out/<tag>/)
# steps: tests | tests:<pytest args> | bench:<args> | ref | ab:<variants,comma-separated>[:bench args] | build:<ntris> | ncu:<workload>
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.used --format=csv > $out/smi.txt 2>&1
nproc > $out/nproc.txt
for step in "$@"; do
  kind=${step%%:*}; arg=${step#*:}; [ "$arg" = "$step" ] && arg=""
  case $kind in
    tests)
      timeout 1500 python -m pytest tests -m gpu -q $arg > $out/pytest.log 2>&1; echo "pytest rc $?" >> $out/pytest.log; tail -5 $out/pytest.log ;;
    testsv)   # tests against a variant library: testsv:<variant>:<pytest args>
      v=${arg%%:*}; pa=${arg#*:}
      SPRAYGPU_LIB=spray_b200/variants/libspraygpu_$v.so timeout 900 python -m pytest -m gpu -x -q $pa > $out/pytest_$v.log 2>&1; echo "pytest rc $?" >> $out/pytest_$v.log; tail -3 $out/pytest_$v.log ;;
    bench)
      n=$(echo "$arg" | tr -c 'a-zA-Z0-9' '_')
      timeout 900 python bench.py $arg > $out/bench_$n.json 2> $out/bench_$n.err; echo "bench $arg rc $?"; tail -c 600 $out/bench_$n.json ;;
    ref)
      timeout 600 python bench.py --impl reference --steps 3 --warmup 1 $arg > $out/bench_ref.json 2> $out/bench_ref.err; tail -c 400 $out/bench_ref.json ;;
    ab)
      vs=${arg%%:*}; ba=${arg#*:}; [ "$ba" = "$arg" ] && ba=""
      for v in $(echo $vs | tr ',' ' '); do
        if [ "$v" = default ]; then lib=spray_b200/libspraygpu.so; else lib=spray_b200/variants/libspraygpu_$v.so; fi
        SPRAYGPU_LIB=$lib timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extra-workloads $ba 2> $out/ab_$v.err | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$v [$ba]', 'Mrays/s %.0f'%d['value'], 'ms %.3f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'e2e %.0f'%d['e2e']['value'], 'nodes/ray %.2f tris/ray %.2f'%(r['nodes_per_ray'], r['tris_per_ray']))" >> $out/ab.txt 2>&1
      done; cat $out/ab.txt ;;
    build)
      timeout 300 python tools/build_one.py $arg > $out/build_$arg.log 2>&1 &&
      ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/build_${arg}_launches.csv python tools/build_one.py $arg > $out/build_${arg}_ncu.log 2>&1
      tail -3 $out/build_$arg.log ;;
    ncu)
      wl=$arg
      timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra-workloads --workload $wl > $out/ncu_plain_$wl.log 2>&1 &&
      ncu --set full --clock-control none --import-source on -k regex:render_tile --launch-skip 4 -c 2 -o $out/prof_$wl python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra-workloads --workload $wl > $out/ncu_$wl.log 2>&1
      ls -la $out/prof_$wl.ncu-rep ;;
    insitu1)  # one-rank wavefront in-situ breakdown: plain run, then the ncu launch list of the same command
      timeout 600 python tools/insitu_one.py $arg 3 > $out/insitu1_$arg.log 2>&1 &&
      ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:p2p_|insitu_|eye_rays|intersect_domains|image_flush' -c 600 --csv --log-file $out/insitu1_${arg}_launches.csv python tools/insitu_one.py $arg 1 > $out/insitu1_${arg}_ncu.log 2>&1
      tail -3 $out/insitu1_$arg.log ;;
    fwd1)    # one-rank ray-forwarding tile under ncu: launch list with durations
      timeout 600 python tools/insitu_one.py $arg 3 fwd > $out/fwd1_$arg.log 2>&1 &&
      ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:fwd_|p2p_' -c 400 --csv --log-file $out/fwd1_${arg}_launches.csv python tools/insitu_one.py $arg 1 fwd > $out/fwd1_${arg}_ncu.log 2>&1
      tail -3 $out/fwd1_$arg.log ;;
    py)      # py:<script and args>  (one GPU)
      lg=$out/py_$(echo "$arg" | tr -c 'a-zA-Z0-9' '_' | cut -c1-60)
      timeout 900 python $arg > $lg.log 2> $lg.err; echo "py $arg rc $?"; tail -8 $lg.log ;;
    smoke)
      timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; tail -2 $out/smoke.log ;;
    env)     # env:NAME=VALUE  (exported for the following steps)
      export "$arg" ;;
    multi)   # multi:<nproc>:<script and args>
      np_=${arg%%:*}; cmd=${arg#*:}
      lg=$out/multi_n${np_}_$(echo "$cmd" | tr -c 'a-zA-Z0-9' '_' | cut -c1-60)${SPRAY_INSITU_RR:+_rr$SPRAY_INSITU_RR}
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np_ --master-addr 127.0.0.1 --master-port 29533 $cmd > $lg.log 2> $lg.err; echo "multi $np_ $cmd rc $?"; tail -c 1500 $lg.log; tail -3 $lg.err ;;
  esac
done
