#!/bin/sh
# Regenerates tests/golden/obj/small_tinyobj.txt with the reference's vendored tinyobjloader (needs /root/reference):
#   make -C oracle ref && sh tests/golden/make_obj_golden.sh
# small.obj / small.mtl / tex.ppm are hand-made fixtures (random vertices, seed 3): triangles only, every index form
# (v, v/vt, v/vt/vn, v//vn, negative), five materials (Pr/Pm, map_Kd, Ns only, illum 7 + Ni, d < 1), a missing material.
set -e
cd "$(dirname "$0")/../.."
oracle/_ref/obj_ref tests/golden/obj/small.obj > tests/golden/obj/small_tinyobj.txt
