cd $GRAFT_REPO_ROOT
for mb in 64 96 128 256 6144; do echo "== cs  MB=$mb"; PCB_SCRATCH_MB=$mb python scripts/exp/kbench.py 32768 24; done
for mb in 64 96 128 256 6144; do echo "== nocs MB=$mb"; PCB_SO=$PWD/scratch_dbg/v_nocs.so PCB_SCRATCH_MB=$mb python scripts/exp/kbench.py 32768 24; done
