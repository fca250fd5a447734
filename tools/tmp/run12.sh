cd $GRAFT_REPO_ROOT
python tools/tmp/h2d_probe.py 2>&1 | tail -12
nvidia-smi topo -m 2>&1 | head -8
nproc; lscpu | grep -i -E "numa|socket|model name" | head -8
