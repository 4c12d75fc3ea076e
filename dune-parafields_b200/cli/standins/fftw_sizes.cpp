// standins/fftw_sizes.cpp -- what the CUDA-path binaries link INSTEAD of
// FFTW3 / FFTW3-MPI / MPI.
//
//  * size queries of FFTW-MPI (the reference's traits class calls them,
//    fieldtraits.hh:204,368-379): answered for one rank, i.e. the whole array
//    is local (the FFTW manual's definition of fftw_mpi_local_size* for a
//    communicator of size 1);
//  * every plan / execute / allocation entry point: ABORTS with a message.
//    A binary linked against this file cannot run a CPU transform, which is
//    the proof that the fields it writes come from the CUDA library;
//  * MPI: single-process answers (rank 0 of 1, collectives are copies,
//    messages to oneself go through a queue).
#include <fftw3-mpi.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <vector>

namespace {
[[noreturn]] void
no_cpu_fft(const char* what)
{
  std::fprintf(stderr,
               "%s called: this binary is linked against the sizes-only FFTW "
               "stand-in; CPU transforms do not exist on the CUDA path\n",
               what);
  std::abort();
}
struct Message
{
  int tag;
  std::vector<unsigned char> data;
};
std::deque<Message> g_queue;

ptrdiff_t
product(int rnk, const ptrdiff_t* n)
{
  ptrdiff_t p = 1;
  for (int i = 0; i < rnk; i++)
    p *= n[i];
  return p;
}
} // namespace

extern "C" {

// ---- MPI, one process -------------------------------------------------------
int MPI_Init(int*, char***) { return 0; }
int MPI_Finalize(void) { return 0; }
int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return 0; }
int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return 0; }
int MPI_Barrier(MPI_Comm) { return 0; }
int MPI_Allreduce(const void* in, void* out, int count, MPI_Datatype t, MPI_Op,
                  MPI_Comm)
{
  if (in != out)
    std::memmove(out, in, (size_t)count * (size_t)t);
  return 0;
}
int MPI_Isend(const void* buf, int count, MPI_Datatype t, int, int tag, MPI_Comm,
              MPI_Request* req)
{
  Message m;
  m.tag = tag;
  m.data.assign((const unsigned char*)buf,
                (const unsigned char*)buf + (size_t)count * (size_t)t);
  g_queue.push_back(std::move(m));
  if (req)
    *req = 0;
  return 0;
}
int MPI_Recv(void* buf, int count, MPI_Datatype t, int, int tag, MPI_Comm,
             MPI_Status*)
{
  for (auto it = g_queue.begin(); it != g_queue.end(); ++it)
    if (it->tag == tag) {
      std::memcpy(buf, it->data.data(),
                  std::min(it->data.size(), (size_t)count * (size_t)t));
      g_queue.erase(it);
      return 0;
    }
  std::fprintf(stderr, "MPI stand-in: receive without a matching send\n");
  std::abort();
}
int MPI_Wait(MPI_Request*, MPI_Status*) { return 0; }
int MPI_Waitall(int, MPI_Request*, MPI_Status*) { return 0; }

// ---- FFTW-MPI: sizes only -----------------------------------------------------
void fftw_mpi_init(void) {}
void fftw_mpi_broadcast_wisdom(MPI_Comm) {}
void fftw_mpi_gather_wisdom(MPI_Comm) {}
int fftw_import_wisdom_from_filename(const char*) { return 0; }
int fftw_export_wisdom_to_filename(const char*) { return 0; }

ptrdiff_t
fftw_mpi_local_size(int rnk, const ptrdiff_t* n, MPI_Comm, ptrdiff_t* local_n0,
                    ptrdiff_t* local_0_start)
{
  *local_n0 = n[0];
  *local_0_start = 0;
  return product(rnk, n);
}
ptrdiff_t
fftw_mpi_local_size_1d(ptrdiff_t n0, MPI_Comm, int, unsigned, ptrdiff_t* local_ni,
                       ptrdiff_t* local_i_start, ptrdiff_t* local_no,
                       ptrdiff_t* local_o_start)
{
  *local_ni = *local_no = n0;
  *local_i_start = *local_o_start = 0;
  return n0;
}
ptrdiff_t
fftw_mpi_local_size_transposed(int rnk, const ptrdiff_t* n, MPI_Comm,
                               ptrdiff_t* local_n0, ptrdiff_t* local_0_start,
                               ptrdiff_t* local_n1, ptrdiff_t* local_1_start)
{
  *local_n0 = n[0];
  *local_0_start = 0;
  *local_n1 = rnk > 1 ? n[1] : 1;
  *local_1_start = 0;
  return product(rnk, n);
}
ptrdiff_t
fftw_mpi_local_size_many_transposed(int rnk, const ptrdiff_t* n,
                                    ptrdiff_t howmany, ptrdiff_t, ptrdiff_t,
                                    MPI_Comm, ptrdiff_t* local_n0,
                                    ptrdiff_t* local_0_start, ptrdiff_t* local_n1,
                                    ptrdiff_t* local_1_start)
{
  *local_n0 = n[0];
  *local_0_start = 0;
  *local_n1 = rnk > 1 ? n[1] : 1;
  *local_1_start = 0;
  return product(rnk, n) * howmany;
}

// ---- everything that would transform on the CPU: abort -------------------------
double* fftw_alloc_real(size_t) { no_cpu_fft("fftw_alloc_real"); }
fftw_complex* fftw_alloc_complex(size_t) { no_cpu_fft("fftw_alloc_complex"); }
void fftw_free(void*) { no_cpu_fft("fftw_free"); }
void fftw_execute(const fftw_plan) { no_cpu_fft("fftw_execute"); }
void fftw_destroy_plan(fftw_plan) { no_cpu_fft("fftw_destroy_plan"); }
fftw_plan
fftw_mpi_plan_dft(int, const ptrdiff_t*, fftw_complex*, fftw_complex*, MPI_Comm,
                  int, unsigned)
{
  no_cpu_fft("fftw_mpi_plan_dft");
}
fftw_plan
fftw_mpi_plan_dft_r2c(int, const ptrdiff_t*, double*, fftw_complex*, MPI_Comm,
                      unsigned)
{
  no_cpu_fft("fftw_mpi_plan_dft_r2c");
}
fftw_plan
fftw_mpi_plan_dft_c2r(int, const ptrdiff_t*, fftw_complex*, double*, MPI_Comm,
                      unsigned)
{
  no_cpu_fft("fftw_mpi_plan_dft_c2r");
}
fftw_plan
fftw_mpi_plan_r2r(int, const ptrdiff_t*, double*, double*, MPI_Comm,
                  const fftw_r2r_kind*, unsigned)
{
  no_cpu_fft("fftw_mpi_plan_r2r");
}
fftw_plan
fftw_mpi_plan_many_r2r(int, const ptrdiff_t*, ptrdiff_t, ptrdiff_t, ptrdiff_t,
                       double*, double*, MPI_Comm, const fftw_r2r_kind*, unsigned)
{
  no_cpu_fft("fftw_mpi_plan_many_r2r");
}

} // extern "C"
