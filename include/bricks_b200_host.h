/*
 * bricks_b200_host.h -- flat C window onto the C++ host templates of include/bricks/.
 *
 * The drop-in host API is C++ (Brick / BrickInfo / BrickStorage, BrickDecomp, init_grid,
 * copyToBrick ...: include/bricks/*.h, mirroring the reference's include/brick.h,
 * bricksetup.h and brick-mpi.h).  Harnesses that are not C++ (bench.py, the pytest parity
 * tests) reach the SAME code through these wrappers, instantiated for 3-D 8x8x8 fp64 bricks
 * -- so the tables the GPU path runs on come from the product's own BrickDecomp, never from
 * the oracle.  Host-only: nothing here touches a device.
 */
#ifndef BRICKS_B200_HOST_H
#define BRICKS_B200_HOST_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bricks_decomp bricks_decomp;
struct bricks_ctx;  /* bricks_b200.h */
struct bricks_step;

/* BrickDecomp<3,8,8,8>::g_region (include/brick-mpi.h:183-191) flattened */
typedef struct {
  uint64_t neighbor;
  unsigned skin_st, skin_ed, pos, len, first_pad, last_pad;
} bricks_region;

enum { BRICKS_SKIN_GOOD = 0, BRICKS_SKIN_NORMAL = 1, BRICKS_SKIN_BAD = 2 };

/* BrickDecomp<3,8,8,8>(dims, depth).initialize(skin3d_*)  -- brick-mpi.h:304-460.
 * dims in elements, contiguous axis first. */
int bricks_decomp_create(const unsigned dims[3], unsigned depth, int skin_kind, bricks_decomp **out);
int bricks_decomp_destroy(bricks_decomp *d);
/* gdims = brick-grid extent including ghost bricks, contiguous axis first */
int bricks_decomp_sizes(const bricks_decomp *d, unsigned *nbricks, unsigned sep_pos[3], unsigned gdims[3],
                        int *nregions, int *nskin);
/* any output may be NULL.  grid: gdims[2]*gdims[1]*gdims[0] cells, i fastest; adj: nbricks*27 */
int bricks_decomp_tables(const bricks_decomp *d, unsigned *grid, unsigned *adj, bricks_region *ghost,
                         bricks_region *skin, long *skin_size);
/* populate(cart, decomp, 0, 1, coo) on a periodic procdims grid (brick-mpi.h:730-753);
 * keys/ranks in allneighbors() order, 27 entries */
int bricks_decomp_rank_map(bricks_decomp *d, const int procdims[3], const int coo[3], uint64_t keys[27],
                           int ranks[27]);
/* byte offsets / sizes / peers of BrickDecomp::exchange (brick-mpi.h:466-495) for a storage of
 * `step` elements per brick, using the rank map of the last bricks_decomp_rank_map call;
 * arrays of nregions entries: the arguments of bricks_exchange_create */
int bricks_decomp_exchange_plan(bricks_decomp *d, size_t step, size_t *recv_off, size_t *send_off,
                                size_t *nbytes, int *recv_peer, int *send_peer);

/* the SHIFT exchange (BrickDecomp::multiStageExchangeView, brick-mpi.h:587-681) as the arguments of
 * one bricks_exchange_create per stage: stage = 0..2 (axis stage+1); writes up to `cap` pieces and
 * returns their number in *npieces (call with cap = 0 to size the arrays) */
int bricks_decomp_shift_plan(bricks_decomp *d, size_t step, int stage, int cap, int *npieces,
                             size_t *recv_off, size_t *send_off, size_t *nbytes, int *recv_peer,
                             int *send_peer);

/* BrickDecomp::makeStep (include/bricks/brick-decomp.h): the library's sweep schedule (bricks_step_* of
 * bricks_b200.h) for this decomposition with the communication program of the chosen exchange --
 * 0 = one send/recv pair per region (BrickDecomp::exchange, brick-mpi.h:466-495), 1 = one packed
 * message per peer rank (the role of ExchangeView, brick-mpi.h:96-123), 2 = the SHIFT exchange
 * (brick-mpi.h:587-681) -- using the rank map of the last bricks_decomp_rank_map call.  The one
 * entry of this header that touches a device (it uploads the tables).  The caller owns *step. */
int bricks_decomp_step_create(bricks_decomp *d, struct bricks_ctx *ctx, int rank, int nranks, int exchange_kind,
                              size_t step_elems, struct bricks_step **step, char *what, size_t what_len);

/* init_grid<3> (bricksetup.h:73-90): grid[n0*n1*n2], adj[n0*n1*n2][27] */
int bricks_init_grid3(const long dimlist[3], unsigned *grid, unsigned *adj);
/* copyToBrick<3> (to_brick != 0) / copyFromBrick<3> (bricksetup.h:172-221) on row-major bricks */
int bricks_copy_brick3(int to_brick, const long dimlist[3], const long padding[3], const long ghost[3],
                       double *arr, const unsigned *grid, double *brick_dat, size_t step, const unsigned *adj,
                       unsigned nbricks);
/* compareBrick<3> (brickcompare.h:30-57): *match = 1 when within BRICK_TOLERANCE; *max_rel = the
 * strict measure max |brick-arr| / |arr| */
int bricks_compare_brick3(const long dimlist[3], const long padding[3], const long ghost[3], double *arr,
                          const unsigned *grid, double *brick_dat, size_t step, const unsigned *adj,
                          unsigned nbricks, int *match, double *max_rel);
/* ---- the 4-D decomposition of weak/main-4d.cpp:98-124: BrickDecomp<4, 4,4,4,4>(dims, depth)
 * .initialize(allneighbors minus the centre).  page_align != 0 gives the numbering of the reference's
 * default build (4x4x4x4 fp64 bricks are half a page: two reserved slots, every region padded to whole
 * pages), 0 the dense numbering of -DDECOMP_PAGEUNALIGN (what its CUDA drivers use).  Tables as for the
 * 3-D calls above with 4 axes and 81 adjacency slots / direction sets. */
typedef struct bricks_decomp4d bricks_decomp4d;
int bricks_decomp4d_create(const unsigned dims[4], unsigned depth, int page_align, bricks_decomp4d **out);
int bricks_decomp4d_destroy(bricks_decomp4d *d);
int bricks_decomp4d_sizes(const bricks_decomp4d *d, unsigned *nbricks, unsigned sep_pos[3], unsigned gdims[4],
                          int *nregions, int *nskin);
int bricks_decomp4d_tables(const bricks_decomp4d *d, unsigned *grid, unsigned *adj, bricks_region *ghost,
                           bricks_region *skin, long *skin_size);
int bricks_decomp4d_rank_map(bricks_decomp4d *d, const int procdims[4], const int coo[4], uint64_t keys[81],
                             int ranks[81]);
int bricks_decomp4d_exchange_plan(bricks_decomp4d *d, size_t step, size_t *recv_off, size_t *send_off,
                                  size_t *nbytes, int *recv_peer, int *send_peer);
/* copyToBrick<4> (to_brick != 0) / copyFromBrick<4> (bricksetup.h:172-221) on row-major 4x4x4x4 bricks */
int bricks_copy_brick4(int to_brick, const long dimlist[4], const long padding[4], const long ghost[4],
                       double *arr, const unsigned *grid, double *brick_dat, size_t step, const unsigned *adj,
                       unsigned nbricks);

/* ---- strong scaling: two-level decomposition (strong/args.cpp:36-56,104-113; strong/main.cu:188-395).
 * A dom^3 domain in (dom/sdom)^3 subdomains named by Z-Morton id; rank r owns a contiguous id
 * range; all subdomains share the BrickDecomp `d` (built for sdom^3).  Offsets are bytes inside
 * one subdomain storage of `step` elements per brick. */
typedef struct bricks_strong_plan bricks_strong_plan;
typedef struct {
  unsigned src_sub, dst_sub;      /* local subdomain indices */
  size_t src_off, dst_off, len;   /* neighbour's skin[i] -> my ghost[i] */
} bricks_strong_link;
typedef struct {
  unsigned sub, region;           /* local subdomain index, BrickDecomp region i */
  size_t off, len, buf_off;       /* range inside the storage, position in the packed message */
  unsigned long sender_id;        /* Morton id of the sending subdomain */
} bricks_strong_piece;
int bricks_strong_plan_create(unsigned dom_size, unsigned sdom_size, int nranks, int rank, const bricks_decomp *d,
                              size_t step, bricks_strong_plan **out);
int bricks_strong_plan_destroy(bricks_strong_plan *p);
/* makeStrongStep (include/bricks/strong-plan.h): the library's sweep schedule for the local subdomains of
 * the plan with the communication program of strong/main.cu:397-447 (pack -> one message per peer ->
 * unpack -> intra-GPU links); subdomain s of the exchanged field lives at a_base + s * sub_bytes on the
 * device.  Touches the device (uploads tables).  The caller owns *step. */
int bricks_strong_step_create(const bricks_strong_plan *p, struct bricks_ctx *ctx, int rank, const bricks_decomp *d,
                              void *a_base, size_t sub_bytes, struct bricks_step **step, size_t *sent_bytes);
int bricks_strong_plan_sizes(const bricks_strong_plan *p, unsigned long *first_id, unsigned long *last_id,
                             size_t *nlinks, int *nsend_peers, int *nrecv_peers);
int bricks_strong_plan_links(const bricks_strong_plan *p, bricks_strong_link *out);
/* dir 0 = send, 1 = receive; peers in ascending rank order */
int bricks_strong_plan_peer(const bricks_strong_plan *p, int dir, int idx, int *rank, size_t *bytes, int *npieces);
int bricks_strong_plan_pieces(const bricks_strong_plan *p, int dir, int idx, bricks_strong_piece *out);
/* owner rank / Morton neighbour of a subdomain id (getrank) and its grid coordinates (i,j,k) */
int bricks_strong_owner(const bricks_strong_plan *p, unsigned long id, int *rank);
int bricks_strong_neighbor(const bricks_strong_plan *p, unsigned long id, uint64_t direction_set, unsigned long *nid);
int bricks_strong_coords(const bricks_strong_plan *p, unsigned long id, unsigned coords[3]);

/* MPI_Dims_create / MPI_Cart_rank / MPI_Cart_coords equivalents (weak/args.cpp:105-115) */
int bricks_cart_dims(int nranks, int ndims, int *dims);
int bricks_cart_rank(int ndims, const int *dims, const int *coo, int *rank);
int bricks_cart_coords(int ndims, const int *dims, int rank, int *coo);
/* ZMORT()[c0][c1][c2] and its inverse (include/zmort.h:18-105) */
unsigned long bricks_zmort3(unsigned long c0, unsigned long c1, unsigned long c2);
unsigned long bricks_zmort3_coord(unsigned long id, unsigned long d);
unsigned long bricks_zmort3_set(unsigned long id, unsigned long d, unsigned long p);
/* randomArray(list, seed) of include/bricks/multiarray.h: n doubles from element `start` */
int bricks_uniform_fill(double *out, size_t n, uint64_t seed, uint64_t start);

#ifdef __cplusplus
}
#endif
#endif /* BRICKS_B200_HOST_H */
