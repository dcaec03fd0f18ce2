/* thread-local error text shared by the host (C) and device (CUDA) halves of librdyb200 */
#ifndef RDYB_ERROR_H
#define RDYB_ERROR_H
#ifdef __cplusplus
extern "C" {
#endif
/* stores a printf-formatted message for rdyb_last_error() and returns `code` */
int rdyb_fail(int code, const char *fmt, ...);
#ifdef __cplusplus
}
#endif

/* PetscCheck-like guard (include/private: PetscCheck(cond, comm, PETSC_ERR_USER, fmt, ...)) */
#define RDYB_CHECK(cond, code, ...)                      \
  do {                                                   \
    if (!(cond)) return rdyb_fail((code), __VA_ARGS__);  \
  } while (0)
/* PetscCall-like propagation */
#define RDYB_CALL(expr)            \
  do {                             \
    int rdyb_ierr_ = (expr);       \
    if (rdyb_ierr_) return rdyb_ierr_; \
  } while (0)
#endif
