# A/B of the chunked upload's split (tools/prof_e2e.py, config 3, CAT_MPC_STREAMED=0) against the streamed form
export CAT_MPC_STREAMED=0
for r in 1 2; do
CAT_MPC_CHUNK_N=7 CAT_MPC_CHUNK_LAST=8192 python tools/prof_e2e.py
python tools/prof_e2e.py
CAT_MPC_CHUNK_N=2 python tools/prof_e2e.py
CAT_MPC_CHUNK_N=3 python tools/prof_e2e.py
CAT_MPC_CHUNK_LAST=9472 python tools/prof_e2e.py
CAT_MPC_CHUNK_LAST=37888 python tools/prof_e2e.py
CAT_MPC_CHUNK_N=2 CAT_MPC_CHUNK_LAST=9472 python tools/prof_e2e.py
CAT_MPC_STREAMED=1 python tools/prof_e2e.py
done
unset CAT_MPC_STREAMED
python tools/prof_e2e_layouts.py
