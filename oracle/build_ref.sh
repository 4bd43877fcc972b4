#!/usr/bin/env bash
# Build the UNMODIFIED uvaia reference (src/nearest.c -> `uvaia`) straight from the sources
# under /root/reference into oracle/_ref/ (git-ignored).  Test infrastructure only: the product
# never links or executes anything from here.  No reference source is copied into this repo.
# Also builds oracle/_ref/libuvaia_ref_score.so exposing the reference's own scoring functions
# (biomcmc_pairwise_score_matches_truncated_idx etc.) for unit-level golden generation.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
R="${UVAIA_REFERENCE:-/root/reference}"
L="$R/submodules/biomcmc-lib/lib"
OUT="$HERE/_ref"
if [ ! -d "$R/src" ]; then echo "reference sources not present at $R (expected on the GPU box): keeping prebuilt $OUT"; exit 0; fi
mkdir -p "$OUT"
LIBSRC=""
for f in hashtable lowlevel random_number_gen constant_random_lists random_number nexus_common bipartition \
  prob_distribution empirical_frequency argtable3 distance_matrix alignment topology_common parsimony genetree \
  reconciliation splitset_distances read_newick_trees char_vector upgma topology_randomise newick_space \
  topology_space topology_distance kmerhash hashfunctions distance_generator clustering_goptics \
  quickselect_quantile fortune_cookies suffix_tree phylogeny likelihood gff3_format file_compression; do
  LIBSRC="$LIBSRC $L/$f.c"
done
CFLAGS="-std=gnu11 -O3 -funroll-loops -fopenmp -w -Dsave_neigbours_to_gz_file=save_neighbours_to_gz_file -I$HERE/cfg -I$HERE/stub -I$L -I$R/src"
gcc $CFLAGS -o "$OUT/uvaia_ref" "$R/src/nearest.c" "$R/src/fastaseq.c" "$R/src/min_heap.c" "$R/src/utils.c" $LIBSRC -lz -lm
# uvaiaball (src/ball.c): the reference's radius search, for the goldens of the uvg_ball_* entry points
gcc $CFLAGS -o "$OUT/uvaiaball_ref" "$R/src/ball.c" "$R/src/fastaseq.c" "$R/src/utils.c" $LIBSRC -lz -lm
gcc $CFLAGS -fPIC -shared -o "$OUT/libuvaia_ref_score.so" "$R/src/fastaseq.c" "$R/src/min_heap.c" "$R/src/utils.c" $LIBSRC -lz -lm
echo "built $OUT/uvaia_ref, $OUT/uvaiaball_ref and $OUT/libuvaia_ref_score.so"
