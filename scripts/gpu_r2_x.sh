# round 2, GPU call X (one GPU): the XML tokenizer tests after the XML-declaration version rule was fixed
set -x
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_legacy.py -m gpu -q -k "xml" > gpurun_out/r2x_tests_xml.log 2>&1; tail -n 3 gpurun_out/r2x_tests_xml.log
