# regenerates profiles/r2_sass_tcgen05.txt from the built objects (see that file's header)
make -C vqvae_b200/csrc -j8 && cd vqvae_b200/csrc/build && for o in knn_tc.o jvp_gemm.o; do echo "== $o"; cuobjdump -sass $o | grep -E "UTCHMMA|LDTM|UBLKCP|UTCBAR" | awk '{print $2}' | sort | uniq -c; done
