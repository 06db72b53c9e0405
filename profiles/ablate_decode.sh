#!/bin/bash
# In-graph cost of each kernel class of the decode step, by leaving it out (B200LLM_SKIP; the switch
# exists only in the tune build: python <pkg>/build.py --tune). profiles/decode_classes.py prints the same as JSON.
export B200LLM_LIB=${B200LLM_LIB:-$(dirname "$0")/../distributed-real-time-chat-and-collaboration-tool_b200/libb200llm_tune.so}
for m in 0 1 2 4 8 16 32 63; do
  echo -n "skip_mask=$m  "; B200LLM_SKIP=$m python profiles/time_decode.py "$@" 2>&1 | grep "graphs=True  pdl=True"
done
