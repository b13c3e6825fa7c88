#!/bin/bash
# One GPU-box pass: parity suite, bench (both arms), ncu launch list, one full ncu capture.
# usage (from the repo root, under gpurun):  bash tools/gpu_round.sh <tag> [skip-ncu]
# Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
set -u
tag=${1:-run}
out=gpurun_out
mkdir -p $out
python -c 'import torch; print(torch.cuda.get_device_name(0))' > $out/${tag}_env.log 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $out/${tag}_env.log 2>&1

timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1
echo "pytest exit $?" >> $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log

timeout 600 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err
echo "bench exit $?"; tail -c 600 $out/${tag}_bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_reference.json 2>> $out/${tag}_bench.err
echo "reference arm exit $?"

if [[ "${2:-}" != "skip-ncu" ]]; then
  small="--steps 2 --warmup 1 --no-extras --no-cpu-baseline --e2e-frames 2000"
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file $out/${tag}_launches.csv python bench.py $small > $out/${tag}_ncu_launches.log 2>&1
  echo "ncu launch list exit $?"
  timeout 900 ncu --set full --clock-control none --import-source on \
      -k 'regex:k_site_states|k_events_compact|k_jumps_spec' -c 3 -f -o $out/${tag}_k1_k2 \
      python bench.py $small > $out/${tag}_ncu_full.log 2>&1
  echo "ncu full exit $?"
  # N4: the RDF kernel on its own (tools/rdf_time.py: diagonal and triclinic cell, both binnings)
  timeout 120 python tools/rdf_time.py > $out/${tag}_rdf_time.txt 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:k_rdf_hist' -c 4 -f \
      -o $out/${tag}_rdf python tools/rdf_time.py > $out/${tag}_ncu_rdf.log 2>&1
  echo "ncu rdf exit $?"
fi
ls -la $out | tail -20
