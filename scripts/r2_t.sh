#!/bin/bash
timeout 600 python -m pytest tests/test_dropout_gpu.py -q -m gpu -k "joint or backward_uses" 2>&1 | grep -E "^E |passed|failed|Error" | head -20
