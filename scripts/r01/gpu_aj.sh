#!/bin/bash
# (E) single-GPU outer look-ahead with the v12 kernels, (F) e2e vs outer block
mkdir -p gpurun_out
run() { tag=$1; shift; w=$1; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline $EXTRA > gpurun_out/aj_${w}_$tag.json 2> gpurun_out/aj_${w}_$tag.err
  echo "== $w $tag"; grep '^{' gpurun_out/aj_${w}_$tag.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); e=d.get('e2e') or {}; print(round(d['value'],1), round(d['ms_per_step'],2), 'e2e', round(e.get('value',0),1), round(e.get('ms_per_step',0),2))"; }
EXTRA="--no-e2e"
run base gauss32768 X=1
run ov2 gauss32768 GAUSS_CUDA_OVERLAP=1 GAUSS_CUDA_OVERLAP_TILES=2
run ov4 gauss32768 GAUSS_CUDA_OVERLAP=1 GAUSS_CUDA_OVERLAP_TILES=4
run ov8 gauss32768 GAUSS_CUDA_OVERLAP=1 GAUSS_CUDA_OVERLAP_TILES=8
run base gauss8192 X=1
run ov4 gauss8192 GAUSS_CUDA_OVERLAP=1 GAUSS_CUDA_OVERLAP_TILES=4
EXTRA="--nb 512"
run e2e_nb512 gauss32768 X=1
EXTRA=""
run e2e_nb1024 gauss32768 X=1
run e2e_nb1024_c8 gauss32768 GAUSS_CUDA_STREAM_CHUNKS=8
