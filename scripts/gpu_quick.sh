#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_contract_score.py -x -q -m gpu -k "row_slabs" 2>&1 | tail -12
