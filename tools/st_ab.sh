# A/B of the streamed upload's row-group split against the chunked upload (tools/prof_e2e.py, config 3)
for r in 1 2; do
CAT_MPC_STREAMED=0 python tools/prof_e2e.py
python tools/prof_e2e.py
CAT_MPC_STREAM_SPLIT=49,1 python tools/prof_e2e.py
CAT_MPC_STREAM_SPLIT=47,3 python tools/prof_e2e.py
CAT_MPC_STREAM_SPLIT=40,8,2 python tools/prof_e2e.py
CAT_MPC_STREAM_SPLIT=45,4,1 python tools/prof_e2e.py
CAT_MPC_STREAM_SPLIT=30,14,5,1 python tools/prof_e2e.py
CAT_MPC_STREAM_SPLIT=24,24,1,1 python tools/prof_e2e.py
CAT_MPC_STREAM_SPLIT=50 python tools/prof_e2e.py
done
