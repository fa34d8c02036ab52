#!/bin/bash
timeout 1200 python -m pytest tests/test_gpu_contract_score.py tests/test_gpu_configs.py tests/test_gpu_qsplit.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5
