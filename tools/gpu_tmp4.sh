set -x
timeout 1200 compute-sanitizer --tool memcheck --print-limit 5 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q -k "reference_chunk_single or malformed or special or capacity or async or two_phase or cfg1 or lookback" 2>&1 | tail -15
timeout 900 compute-sanitizer --tool racecheck --print-limit 5 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lookback or two_phase" 2>&1 | tail -12
