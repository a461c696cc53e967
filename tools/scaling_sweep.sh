#!/bin/bash
# One 8-GPU box call: strong-scaling sweep with the device timeline on, results under gpurun_out/.
#   /usr/local/graft/bin/gpurun --gpus 8 --timeout 400 -- 'bash tools/scaling_sweep.sh "8 4 2 1" [extra bench flags]'
# then here:  python tools/summarise_trace.py gpurun_out/scale_n*.err
# variants: prefix the command with EVP_B200_LOOP=0 (a launch per subcycle; the device timeline needs it) or
# EVP_B200_LIB=cice5_b200/lib/exp_<name>.so (python -m cice5_b200.build --variant <name> -D...)
NS=${1:-"8 4 2 1"}
shift
export EVP_B200_TRACE=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
mkdir -p gpurun_out
for n in $NS; do
  if [ "$n" = "1" ]; then
    timeout 200 python bench.py --no-cpu-baseline --no-e2e "$@" > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else
    timeout 200 $TR --nproc-per-node $n --master-port $((29500 + n)) bench.py --gpus $n --no-cpu-baseline --no-e2e "$@" \
      > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  grep '^{' gpurun_out/scale_n$n.json | cut -c1-140
done
