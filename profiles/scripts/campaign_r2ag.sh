#!/bin/bash
set -u
O=gpurun_out/r2al
mkdir -p $O
python profiles/scripts/attn_bench.py 20 cfg5 long > $O/attn.txt 2>&1; cat $O/attn.txt
python profiles/scripts/with_lib.py ftop profiles/scripts/attn_bench.py 20 cfg5 long > $O/attn_ftop.txt 2>&1; cat $O/attn_ftop.txt
