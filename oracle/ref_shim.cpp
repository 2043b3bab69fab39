// ref_shim.cpp -- extern "C" doorway onto the UNMODIFIED reference loader
// (/root/reference/ReadMatrixMarket/loadMatrixMarket.cpp:47-253), which is compiled from
// where it lies into oracle/_ref/libloadmatrix_ref.so by oracle/Makefile.
// TEST INFRASTRUCTURE ONLY: used to pin our own loader (respasol_b200/host/rsg_loader.cpp)
// and to generate tests/golden/loader_*.json.  No reference source is copied here.
#include "loadMatrixMarket.h"

extern "C" {

// returns the reference's own return value (1 = success, 0 = failure)
int ref_loadMatrixMarket(const char *file, CSR *out, int outputBase, int transpose)
{
    return loadMatrixMarket(file, out, outputBase, transpose);
}

void ref_free_csr(CSR *a)
{
    FREE(a->rowptr);
    FREE(a->colidx);
    FREE(a->values);
}

// the reference's per-row sort (loadMatrixMarket.cpp:5-26), exposed so the
// duplicate-column ordering of our loader can be pinned against it
void ref_qsort(int *idx, double *w, int left, int right) { qsort(idx, w, left, right); }

}
