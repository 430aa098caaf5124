#!/bin/bash
# Every differential fuzzer of the repository, CPU arms (a few minutes each on 8 cores).  The --gpu arms of fuzz_traversal.py and
# fuzz_scatter.py need a B200 (tools/r02_fuzz_gpu.sh, tools/r02_fuzz_scatter_gpu.sh).  Needs /root/reference for the arms that
# compile the reference's own code (oracle/_ref/libref.so, libref_host.so); without it only the first two run.
set -e
cd "$(dirname "$0")/.."
S=${1:-1}
python tools/fuzz_traversal.py --scenes 400 --rays 1000 --seed $S          # product traversal vs brute force / the oracle's traversal
python tools/fuzz_scatter.py --scenes 300 --seed $S                         # product scatter vs the oracle (exact)
if [ -d /root/reference ] || [ -f oracle/_ref/libref_host.so ]; then
  python tools/fuzz_scatter.py --scenes 200 --seed $S --width 48 --height 32 --reference   # + the reference's own shaders
  python tools/fuzz_pack.py --scenes 150 --seed $S                          # PackSceneData incl. edit rounds vs the reference's
  python tools/fuzz_obj.py --files 300 --seed $S                            # OBJ import vs LoadModelAsPrefab + tinyobjloader
  python tools/fuzz_image_files.py --files 1500 --seed $S                   # image decoders vs stb_image on valid random files
fi
echo "all fuzzers: no difference"
