/* standins/fftw3.h -- declarations of the FFTW3 names the reference's headers
 * mention (backends/fftwwrapper.hh:230-415), so that they COMPILE next to the
 * CUDA path.  No transform is implemented: see fftw_sizes.cpp. */
#ifndef PF_STANDIN_FFTW3_H
#define PF_STANDIN_FFTW3_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef double fftw_complex[2];
typedef struct pf_no_plan* fftw_plan;
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
