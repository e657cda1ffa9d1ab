./tests/cpp/test_compressed 2>&1 | tail -8
(timeout 600 python -m pytest tests/test_cpp_host.py tests/test_abi.py -m gpu -x -q 2>&1 | tail -3)
