/* forcing_ref_stubs.c — link-time stand-ins for the PETSc / RDy entry points that the reference's rdyforcing_dataset.c and
 * rdyforcing_map.c reference on paths forcing_ref_driver.c never takes (file viewers, centroid getters).  TEST INFRASTRUCTURE ONLY.
 * Kept apart from the driver so that no reference header (with the real prototypes) is in scope. */
#include <stdio.h>
#define STUB(name) int name(void) { fprintf(stderr, "forcing oracle: %s is not available in the shim\n", #name); return 83; }
STUB(PetscViewerBinaryOpen) STUB(PetscViewerDestroy) STUB(VecCreate) STUB(VecCreateSeq) STUB(VecLoad)
STUB(RDyCreateOneDOFGlobalVec) STUB(RDyReadOneDOFGlobalVecFromBinaryFile) STUB(RDyWriteOneDOFGlobalVecToBinaryFile)
STUB(RDyGetOwnedCellXCentroids) STUB(RDyGetOwnedCellYCentroids) STUB(RDyGetBoundaryEdgeXCentroids) STUB(RDyGetBoundaryEdgeYCentroids)
int VecDestroy(void *v) { (void)v; return 0; }
