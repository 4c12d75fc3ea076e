/* oracle/shim/fftw3.h -- declarations of the FFTW3 (double) entry points the
 * reference calls (backends/fftwwrapper.hh:230-415).  TEST INFRASTRUCTURE:
 * FFTW3 / FFTW3-MPI are third-party libraries absent from this image (version
 * unpinned by the reference); oracle/shim/fftw_shim.cpp implements their
 * published transform definitions (FFTW manual, "What FFTW really computes")
 * with a plain FFT so that the reference's own loops can run single-rank. */
#ifndef PF_SHIM_FFTW3_H
#define PF_SHIM_FFTW3_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef double fftw_complex[2];
typedef struct pf_shim_plan* fftw_plan;
typedef enum
{
  FFTW_R2HC = 0, FFTW_HC2R = 1, FFTW_DHT = 2, FFTW_REDFT00 = 3,
  FFTW_REDFT01 = 4, FFTW_REDFT10 = 5, FFTW_REDFT11 = 6, FFTW_RODFT00 = 7,
  FFTW_RODFT01 = 8, FFTW_RODFT10 = 9, FFTW_RODFT11 = 10
} fftw_r2r_kind;

#define FFTW_FORWARD (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_MEASURE (0U)
#define FFTW_ESTIMATE (1U << 6)

double* fftw_alloc_real(size_t n);
fftw_complex* fftw_alloc_complex(size_t n);
void fftw_free(void* p);
void fftw_execute(const fftw_plan p);
void fftw_destroy_plan(fftw_plan p);
int fftw_import_wisdom_from_filename(const char*);
int fftw_export_wisdom_to_filename(const char*);

#ifdef __cplusplus
}
#endif
#endif
