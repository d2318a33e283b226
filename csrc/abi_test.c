/*
 * C driver for the host half of the ABI in include/fmpi.h — the same entry points the Julia
 * binding reaches with ccall (julia/FluxMPIB200/src/host.jl), exercised from plain C so the
 * contract is tested on a machine that has no Julia. It forks W ranks, joins the shared-memory
 * world and checks allreduce / bcast / reduce (non-root untouched) / allgather / barrier, i.e.
 * the behaviour /root/reference/test/test_mpi_extensions.jl:34-62 asserts for MPI.
 *
 *   gcc -std=c11 -O2 -Iinclude csrc/abi_test.c -Lfluxmpi_b200/_lib -lfmpi_host -o abi_test
 *   LD_LIBRARY_PATH=fluxmpi_b200/_lib ./abi_test 3
 */
#define _POSIX_C_SOURCE 200809L
#define FMPI_HOST_ONLY
#include "fmpi.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/wait.h>
#include <unistd.h>

#define CHECK(call)                                                                  \
  do {                                                                               \
    if ((call) != 0) {                                                               \
      fprintf(stderr, "rank %d: %s failed: %s\n", rank, #call, fmpi_last_error());   \
      return 1;                                                                      \
    }                                                                                \
  } while (0)

static int run_rank(const char* name, int rank, int world, int nodes, int port) {
  fmpi_host_comm* c = NULL;
  if (nodes == 1) {
    CHECK(fmpi_host_init(name, rank, world, 1 << 16, 30.0, &c));
  } else { /* node-major ranks: shared memory per node, TCP mesh between the node leaders */
    const int per = world / nodes;
    char node_name[96];
    snprintf(node_name, sizeof(node_name), "%s_n%d", name, rank / per);
    CHECK(fmpi_host_init_multi(node_name, rank % per, per, rank / per, nodes, "127.0.0.1", port,
                               1 << 16, 30.0, &c));
  }
  if (fmpi_host_rank(c) != rank || fmpi_host_world(c) != world) return 2;

  /* allreduce, in place, a message larger than one staging slot (chunking) */
  enum { N = 50000 };
  double* x = malloc(sizeof(double) * N);
  for (int i = 0; i < N; ++i) x[i] = (double)(rank + 1);
  CHECK(fmpi_host_allreduce(c, x, x, N, FMPI_F64, FMPI_SUM));
  for (int i = 0; i < N; ++i)
    if (x[i] != world * (world + 1) / 2.0) { fprintf(stderr, "rank %d: allreduce sum wrong\n", rank); return 3; }
  for (int i = 0; i < N; ++i) x[i] = 1.0;
  CHECK(fmpi_host_allreduce(c, x, x, N, FMPI_F64, FMPI_PROD));
  if (x[N - 1] != 1.0) return 4;

  /* bcast: root ones, others zeros -> ones everywhere */
  float b[6];
  for (int i = 0; i < 6; ++i) b[i] = rank == 0 ? 1.0f : 0.0f;
  CHECK(fmpi_host_bcast(c, b, sizeof(b), 0));
  for (int i = 0; i < 6; ++i)
    if (b[i] != 1.0f) return 5;

  /* reduce to the last rank: non-root buffers stay as they were */
  int32_t r4[4] = {rank, rank, rank, rank}, out[4] = {-7, -7, -7, -7};
  const int root = world - 1;
  CHECK(fmpi_host_reduce(c, r4, out, 4, FMPI_I32, FMPI_MAX, root));
  for (int i = 0; i < 4; ++i)
    if (out[i] != (rank == root ? world - 1 : -7)) { fprintf(stderr, "rank %d: reduce wrong\n", rank); return 6; }

  /* allgather of one word per rank */
  int32_t mine = 100 + rank;
  int32_t* all = malloc(sizeof(int32_t) * world);
  CHECK(fmpi_host_allgather(c, &mine, all, sizeof(mine)));
  for (int r = 0; r < world; ++r)
    if (all[r] != 100 + r) return 7;

  /* three arrays as one message (pieces straddle the 64 KB slots) */
  enum { NA = 5, NB = 30000, NC = 1 };
  double *ta = malloc(sizeof(double) * NA), *tb = malloc(sizeof(double) * NB), tc[NC] = {rank + 1.0};
  for (int i = 0; i < NA; ++i) ta[i] = rank + 1.0;
  for (int i = 0; i < NB; ++i) tb[i] = (rank + 1.0) * (i % 7);
  void* leaves[3] = {ta, tb, tc};
  const uint64_t counts[3] = {NA, NB, NC};
  CHECK(fmpi_host_allreduce_many(c, 3, leaves, counts, FMPI_F64, FMPI_SUM));
  const double tri = world * (world + 1) / 2.0;
  for (int i = 0; i < NB; ++i)
    if (tb[i] != tri * (i % 7)) { fprintf(stderr, "rank %d: allreduce_many wrong at %d\n", rank, i); return 8; }
  if (ta[NA - 1] != tri || tc[0] != tri) return 9;

  /* symmetric host memory: all-reduce in place on a range that starts inside the block */
  fmpi_host_symm* seg = NULL;
  CHECK(fmpi_host_symm_alloc(c, sizeof(float) * 100000, &seg));
  float* sm = (float*)fmpi_host_symm_ptr(seg);
  if (!sm || fmpi_host_symm_bytes(seg) < sizeof(float) * 100000) return 12;
  for (int i = 0; i < 100000; ++i) sm[i] = (float)(rank + 1);
  CHECK(fmpi_host_allreduce_symm(c, seg, sizeof(float) * 7, 90001, FMPI_F32, FMPI_SUM));
  for (int i = 0; i < 100000; ++i) {
    const float want = (i >= 7 && i < 7 + 90001) ? (float)tri : (float)(rank + 1);
    if (sm[i] != want) { fprintf(stderr, "rank %d: allreduce_symm wrong at %d\n", rank, i); return 13; }
  }
  if (fmpi_host_allreduce_symm(c, seg, 0, 1u << 30, FMPI_F32, FMPI_SUM) == 0) return 14; /* out of range */
  CHECK(fmpi_host_symm_free(c, seg));

  /* the rule loop: x -= eta * g over two leaves */
  float p1[3] = {1, 1, 1}, p2[1] = {4}, g1[3] = {2, 2, 2}, g2[1] = {8};
  void* ps[2] = {p1, p2};
  const void* gs[2] = {g1, g2};
  const uint64_t ns[2] = {3, 1};
  const double hyper[7] = {0.5, 0.9, 0.999, 1e-8, 0.0, 0.0, 1.0};
  CHECK(fmpi_host_multi_tensor_step(2, ps, gs, NULL, NULL, ns, FMPI_F32, FMPI_RULE_DESCENT, hyper, 0));
  if (p1[2] != 0.0f || p2[0] != 0.0f) return 10;
  if (fmpi_host_multi_tensor_step(2, ps, gs, NULL, NULL, ns, FMPI_F32, FMPI_RULE_ADAM, hyper, 0) == 0)
    return 11; /* Adam without moment arrays must be refused */

  CHECK(fmpi_host_barrier(c));
  CHECK(fmpi_host_finalize(c));
  free(ta);
  free(tb);
  free(x);
  free(all);
  return 0;
}

int main(int argc, char** argv) {
  const int world = argc > 1 ? atoi(argv[1]) : 3;
  const int nodes = argc > 2 ? atoi(argv[2]) : 1; /* > 1: pretend nodes on this machine */
  const int port = argc > 3 ? atoi(argv[3]) : 0;
  char name[64];
  snprintf(name, sizeof(name), "/fmpi_abitest_%d", (int)getpid());
  pid_t kids[FMPI_MAX_RANKS];
  if (world < 1 || world > FMPI_MAX_RANKS || nodes < 1 || world % nodes) return 2;
  for (int r = 0; r < world; ++r) {
    kids[r] = fork();
    if (kids[r] == 0) _exit(run_rank(name, r, world, nodes, port));
  }
  int bad = 0;
  for (int r = 0; r < world; ++r) {
    int st = 0;
    waitpid(kids[r], &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) {
      fprintf(stderr, "rank %d exited with %d\n", r, WIFEXITED(st) ? WEXITSTATUS(st) : -1);
      bad = 1;
    }
  }
  if (!bad) printf("ABI_OK world=%d\n", world);
  return bad;
}
