/*
 * oracle/refshim/mpi_stub/mpi.h -- TEST INFRASTRUCTURE, not product code.
 *
 * A single-process stand-in for <mpi.h>, just large enough that the reference's
 * include/brick-mpi.h and stencils/fake.h compile when built by oracle/Makefile
 * into oracle/_ref/.  No MPI library exists in this image.  Only the calls the
 * reference headers name are declared; every data-moving call aborts (the shim
 * never reaches them), and MPI_Cart_rank is a row-major periodic Cartesian rank
 * (the MPI-standard ordering, which the reference relies on at
 * include/brick-mpi.h:734) over the process grid set with refstub_set_cart().
 */
#ifndef BRICKS_B200_ORACLE_MPI_STUB_H
#define BRICKS_B200_ORACLE_MPI_STUB_H

#include <cstdio>
#include <cstdlib>

typedef int MPI_Comm;
typedef int MPI_Request;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Win;
typedef struct { int unused; } MPI_Status;

#define MPI_COMM_NULL (-1)
#define MPI_COMM_WORLD 0
#define MPI_CHAR 1
#define MPI_DOUBLE 2
#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_LAND 4
#define MPI_THREAD_SERIALIZED 2

extern "C" {
extern int refstub_cart_ndims;
extern int refstub_cart_dims[8];
}

static inline void refstub_set_cart(int nd, const int *dims) {
  refstub_cart_ndims = nd;
  for (int i = 0; i < nd; ++i) refstub_cart_dims[i] = dims[i];
}

static inline int refstub_unreachable(const char *what) {
  std::fprintf(stderr, "oracle mpi stub: %s must not be called\n", what);
  std::abort();
  return 1;
}

static inline int MPI_Cart_rank(MPI_Comm, const int *coords, int *rank) {
  int r = 0;
  for (int d = 0; d < refstub_cart_ndims; ++d) {
    int n = refstub_cart_dims[d];
    int c = ((coords[d] % n) + n) % n;
    r = r * n + c;
  }
  *rank = r;
  return 0;
}

static inline int MPI_Comm_size(MPI_Comm, int *size) {
  int s = 1;
  for (int d = 0; d < refstub_cart_ndims; ++d) s *= refstub_cart_dims[d];
  *size = s;
  return 0;
}
static inline int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = 0; return 0; }
static inline int MPI_Barrier(MPI_Comm) { return 0; }
static inline int MPI_Irecv(void *, size_t, MPI_Datatype, int, int, MPI_Comm, MPI_Request *) {
  return refstub_unreachable("MPI_Irecv");
}
static inline int MPI_Isend(const void *, size_t, MPI_Datatype, int, int, MPI_Comm, MPI_Request *) {
  return refstub_unreachable("MPI_Isend");
}
static inline int MPI_Waitall(int, MPI_Request *, MPI_Status *) {
  return refstub_unreachable("MPI_Waitall");
}
static inline int MPI_Reduce(const void *, void *, int, MPI_Datatype, MPI_Op, int, MPI_Comm) {
  return refstub_unreachable("MPI_Reduce");
}
static inline int MPI_Win_fence(int, MPI_Win) { return refstub_unreachable("MPI_Win_fence"); }
static inline int MPI_Get(void *, size_t, MPI_Datatype, int, size_t, size_t, MPI_Datatype, MPI_Win) {
  return refstub_unreachable("MPI_Get");
}

#endif
