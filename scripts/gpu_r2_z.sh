# round 2, GPU call Z (one GPU): smoke() on the last build of the round
set -x
mkdir -p gpurun_out
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.log 2>&1; tail -n 1 gpurun_out/r2z_smoke.log | cut -c1-200
