# round 2, GPU call M (one GPU): keys-kernel thread-count / register variants, finer sweep of the host call's chunking
set -x
mkdir -p gpurun_out
timeout 900 python scripts/time_variants.py > gpurun_out/r2m_variants.json 2> gpurun_out/r2m_variants.err; tail -n 12 gpurun_out/r2m_variants.err | cut -c1-250
timeout 900 python scripts/sweep_host_pipeline.py > gpurun_out/r2m_host_pipeline.json 2> gpurun_out/r2m_host_pipeline.err; tail -n 12 gpurun_out/r2m_host_pipeline.err | cut -c1-300
