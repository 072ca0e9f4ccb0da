/* oracle/ccicp_port.h -- TEST INFRASTRUCTURE ONLY: C ABI of the CPU restatement
 * (ccicp_port.c).  Same entry points and argument meaning as the ccref_* probes
 * around the unmodified reference (oracle/ref_probe.cpp); oracle/refbind.py
 * binds both. */
#ifndef CCICP_PORT_H
#define CCICP_PORT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
enum { CCPORT_OK = 0, CCPORT_TOO_FEW_MODEL = 1, CCPORT_TOO_FEW_SCENE = 2, CCPORT_NO_CORR = 4 };

int ccport_fit(const double* mGA, int nmGA, const double* mNGA, int nmNGA, const double* sGA, int nsGA,
               const double* sNGA, int nsNGA, int max_iter, double min_delta, double indist, double* Rt,
               int* num_corr, double* eW);
int ccport_trace(const double* mGA, int nmGA, const double* mNGA, int nmNGA, const double* sGA, int nsGA,
                 const double* sNGA, int nsNGA, int max_iter, double min_delta, double indist, const double* Rt0,
                 double* Rt_trace, int32_t* corr, float* dis, uint8_t* tie, int32_t* ncorr, double* delta,
                 int* iters_run, double* eW);
int ccport_tree(const double* pts, int n, int32_t* ind_out, float* nodes_out, int max_nodes);
void ccport_nn(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis);
/* 2x2 SVD of row-major H: U, W(2), V as Matrix::svd leaves them (matrix.cpp:582-810) */
void ccport_svd2(const double* H, double* U, double* W, double* V);
#ifdef __cplusplus
}
#endif
#endif
