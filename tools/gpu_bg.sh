#!/bin/bash
# after the object-bounds field: the bridge's descriptions of the instanced scenes again (fixtures), then every golden re-measured
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/desc
export B200PT_REQUIRE_GPU=1
bash tools/gpu_bc.sh 2>&1 | grep -E 'instances|normals|bridge failed' | head
bash tools/gpu_bd.sh 2>&1 | tail -12
