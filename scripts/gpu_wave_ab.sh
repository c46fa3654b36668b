# usage: bash scripts/gpu_wave_ab.sh <tag> — wavefront A/B over build/var/*.so: whole frames and the one-GPU shard probe
tag=$1
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/${tag}_pytest.log; tail -3 gpurun_out/${tag}_pytest.log
export RTC_WAVEFRONT=1
bash scripts/gpu_variants.sh $tag reflect_refract_1080p cover_1080p teapot_standin_1080p mesh_and_spheres_4k
for so in build/var/*.so; do
  RTC_CUDA_LIB=$PWD/$so timeout 300 python scripts/shard_probe.py 2>&1 | grep "tile_rows=8\|full" | sed "s/^/$(basename $so .so) /"
done
