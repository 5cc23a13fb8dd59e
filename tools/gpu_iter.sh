#!/bin/bash
python -m pytest tests/test_gpu_bert.py tests/test_serving.py tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -2
python tools/probe_bert_small.py 2>&1 | grep "ms per" | cut -c1-44
TUTOR_BERT_NO_GRAPH=1 python tools/probe_bert_small.py 2>&1 | grep "ms per" | cut -c1-44
python tools/probe_bert_e2e.py 2>&1 | tail -4
