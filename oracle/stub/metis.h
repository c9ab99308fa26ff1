/* TEST INFRASTRUCTURE ONLY -- stand-in for <metis.h> (the reference's METIS is an un-vendored submodule,
 * pvc1989/METIS @ 3d8fc59): just the types and prototypes mesh/metis.hpp names, so that the reference's own
 * GetNodeParts / GetNewOrder / ShuffleData / ShuffleConnectivity templates compile.  Nothing here is ever called. */
#ifndef ORACLE_STUB_METIS_H_
#define ORACLE_STUB_METIS_H_
#include <stdint.h>
typedef int64_t idx_t; /* = cgsize_t of the 64-bit CGNS build the reference asserts (mesh/mapper.hpp:28) */
typedef float real_t;
#define METIS_OK 1
#define METIS_NOPTIONS 40
#ifdef __cplusplus
extern "C" {
#endif
int METIS_Free(void *ptr);
int METIS_MeshToDual(idx_t *ne, idx_t *nn, idx_t *eptr, idx_t *eind, idx_t *ncommon, idx_t *numflag, idx_t **r_xadj,
                     idx_t **r_adjncy);
int METIS_PartGraphKway(idx_t *nvtxs, idx_t *ncon, idx_t *xadj, idx_t *adjncy, idx_t *vwgt, idx_t *vsize, idx_t *adjwgt,
                        idx_t *nparts, real_t *tpwgts, real_t *ubvec, idx_t *options, idx_t *edgecut, idx_t *part);
int METIS_PartMeshDual(idx_t *ne, idx_t *nn, idx_t *eptr, idx_t *eind, idx_t *vwgt, idx_t *vsize, idx_t *ncommon,
                       idx_t *nparts, real_t *tpwgts, idx_t *options, idx_t *objval, idx_t *epart, idx_t *npart);
#ifdef __cplusplus
}
#endif
#endif
