cd $GRAFT_REPO_ROOT
bash tools/profile.sh r02g standard 2>&1 | tail -8
