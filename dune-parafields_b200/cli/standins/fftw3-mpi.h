/* standins/fftw3-mpi.h -- the FFTW3-MPI names of the reference's headers.
 * Only the SIZE queries are answered (fieldtraits.hh:204,368-379 asks FFTW-MPI
 * for the slab decomposition); every plan / execute entry point aborts
 * (fftw_sizes.cpp). */
#ifndef PF_STANDIN_FFTW3_MPI_H
#define PF_STANDIN_FFTW3_MPI_H
#include <fftw3.h>
#include <mpi.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FFTW_MPI_TRANSPOSED_IN (1U << 29)
#define FFTW_MPI_TRANSPOSED_OUT (1U << 30)

void fftw_mpi_init(void);
void fftw_mpi_broadcast_wisdom(MPI_Comm);
void fftw_mpi_gather_wisdom(MPI_Comm);

ptrdiff_t fftw_mpi_local_size(int rnk, const ptrdiff_t* n, MPI_Comm comm,
                              ptrdiff_t* local_n0, ptrdiff_t* local_0_start);
ptrdiff_t fftw_mpi_local_size_1d(ptrdiff_t n0, MPI_Comm comm, int sign,
                                 unsigned flags, ptrdiff_t* local_ni,
                                 ptrdiff_t* local_i_start, ptrdiff_t* local_no,
                                 ptrdiff_t* local_o_start);
ptrdiff_t fftw_mpi_local_size_transposed(int rnk, const ptrdiff_t* n,
                                         MPI_Comm comm, ptrdiff_t* local_n0,
                                         ptrdiff_t* local_0_start,
                                         ptrdiff_t* local_n1,
                                         ptrdiff_t* local_1_start);
ptrdiff_t fftw_mpi_local_size_many_transposed(
  int rnk, const ptrdiff_t* n, ptrdiff_t howmany, ptrdiff_t block0,
  ptrdiff_t block1, MPI_Comm comm, ptrdiff_t* local_n0,
  ptrdiff_t* local_0_start, ptrdiff_t* local_n1, ptrdiff_t* local_1_start);

fftw_plan fftw_mpi_plan_dft(int rnk, const ptrdiff_t* n, fftw_complex* in,
                            fftw_complex* out, MPI_Comm comm, int sign,
                            unsigned flags);
fftw_plan fftw_mpi_plan_dft_r2c(int rnk, const ptrdiff_t* n, double* in,
                                fftw_complex* out, MPI_Comm comm,
                                unsigned flags);
fftw_plan fftw_mpi_plan_dft_c2r(int rnk, const ptrdiff_t* n, fftw_complex* in,
                                double* out, MPI_Comm comm, unsigned flags);
fftw_plan fftw_mpi_plan_r2r(int rnk, const ptrdiff_t* n, double* in,
                            double* out, MPI_Comm comm,
                            const fftw_r2r_kind* kind, unsigned flags);
fftw_plan fftw_mpi_plan_many_r2r(int rnk, const ptrdiff_t* n, ptrdiff_t howmany,
                                 ptrdiff_t iblock, ptrdiff_t oblock, double* in,
                                 double* out, MPI_Comm comm,
                                 const fftw_r2r_kind* kind, unsigned flags);

#ifdef __cplusplus
}
#endif
#endif
