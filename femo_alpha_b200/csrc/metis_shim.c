/* k-way graph partitioning of the cell-adjacency graph through METIS ("METIS k-way on the cell-adjacency
 * graph for wing boxes -- libmetis_static.a ships with CUDA 12.9, header prototypes declared by hand").  Host only; built
 * into libfemo_metis.so by femo_alpha_b200/build.py when the archive is present.  The reference has no partitioner (it
 * is serial).  The archive is the one cuSOLVER's sparse routines link: METIS 5.x with 64-bit idx_t (probed:
 * METIS_SetDefaultOptions writes 40 x 8 bytes). */
#include <stdint.h>
#include <stdlib.h>

typedef int64_t idx_t;
#define METIS_NOPTIONS 40
#define METIS_OPTION_SEED 8
#define METIS_OPTION_CONTIG 11
#define METIS_OPTION_NUMBERING 17
#define METIS_OK 1

int METIS_SetDefaultOptions(idx_t* options);
int METIS_PartGraphKway(idx_t* nvtxs, idx_t* ncon, idx_t* xadj, idx_t* adjncy, idx_t* vwgt, idx_t* vsize, idx_t* adjwgt,
                        idx_t* nparts, void* tpwgts, void* ubvec, idx_t* options, idx_t* objval, idx_t* part);
int METIS_PartGraphRecursive(idx_t* nvtxs, idx_t* ncon, idx_t* xadj, idx_t* adjncy, idx_t* vwgt, idx_t* vsize, idx_t* adjwgt,
                             idx_t* nparts, void* tpwgts, void* ubvec, idx_t* options, idx_t* objval, idx_t* part);

/* xadj (n + 1), adjncy: CSR of the undirected graph (both directions stored, no self loops), 0-based.
 * part (n) receives the part of every vertex; returns 0 on success, the METIS status otherwise; *edge_cut = cut edges. */
int femo_metis_kway(int64_t n, const int64_t* xadj, const int64_t* adjncy, int64_t nparts, int contiguous, int64_t seed,
                    int64_t* part, int64_t* edge_cut) {
  if (n <= 0 || nparts < 1 || !xadj || !adjncy || !part) return -1;
  if (nparts == 1) { for (int64_t i = 0; i < n; ++i) part[i] = 0; if (edge_cut) *edge_cut = 0; return 0; }
  idx_t options[METIS_NOPTIONS];
  METIS_SetDefaultOptions(options);
  options[METIS_OPTION_NUMBERING] = 0;
  options[METIS_OPTION_SEED] = seed;              /* fixed seed: the same mesh gives the same partition on every rank */
  options[METIS_OPTION_CONTIG] = contiguous ? 1 : 0;
  idx_t nv = n, ncon = 1, np = nparts, cut = 0;
  /* METIS wants non-const arrays */
  int r = METIS_PartGraphKway(&nv, &ncon, (idx_t*)xadj, (idx_t*)adjncy, NULL, NULL, NULL, &np, NULL, NULL, options, &cut, part);
  if (edge_cut) *edge_cut = cut;
  return r == METIS_OK ? 0 : (r == 0 ? -2 : r);
}
