#!/bin/bash
# In-graph cost of each kernel class of the decode step, by leaving it out (B200LLM_SKIP).
for m in 0 1 2 4 8 16 32 63; do
  echo -n "skip_mask=$m  "; B200LLM_SKIP=$m python profiles/time_decode.py "$@" 2>&1 | grep "graphs=True  pdl=True"
done
