// Self-test of the MPI stand-in on host buffers (no GPU needed): collectives, tag matching out
// of posting order, probe + get_count, zero-length messages, "post all, wait in any order".
#include <mpi.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define REQUIRE(c)                                                                                 \
  do {                                                                                             \
    if (!(c)) {                                                                                    \
      std::fprintf(stderr, "rank %d: FAILED %s (line %d)\n", rank, #c, __LINE__);                  \
      MPI_Abort(MPI_COMM_WORLD, 3);                                                                \
    }                                                                                              \
  } while (0)

int main(int argc, char **argv) {
  MPI_Init(&argc, &argv);
  int rank, size;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  // bcast of a large buffer (several collective chunks)
  std::vector<int> big(300000);
  if (rank == 0) for (size_t i = 0; i < big.size(); ++i) big[i] = int(i * 7);
  MPI_Bcast(big.data(), int(big.size()), MPI_INT, 0, MPI_COMM_WORLD);
  for (size_t i = 0; i < big.size(); i += 9973) REQUIRE(big[i] == int(i * 7));
  // allreduce max / sum
  double t[3] = {double(rank), -double(rank), 1.0};
  MPI_Allreduce(MPI_IN_PLACE, t, 3, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);
  REQUIRE(t[0] == size - 1 && t[1] == 0 && t[2] == 1.0);
  int one = 1, total = 0;
  MPI_Allreduce(&one, &total, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD);
  REQUIRE(total == size);
  // ring: everybody posts 3 tagged sends to next and 3 receives from prev, waits in another order
  const int next = (rank + 1) % size, prev = (rank + size - 1) % size;
  std::vector<std::vector<float>> out(3), in(3);
  MPI_Request sr[3], rr[3];
  for (int k = 0; k < 3; ++k) {
    out[k].assign(size_t(1000 * (k + 1)), float(rank * 10 + k));
    in[k].assign(size_t(1000 * (k + 1)), -1.f);
  }
  for (int k = 2; k >= 0; --k) MPI_Irecv(in[k].data(), int(in[k].size()), MPI_FLOAT, prev, 100 + k, MPI_COMM_WORLD, &rr[k]);
  for (int k = 0; k < 3; ++k) MPI_Isend(out[k].data(), int(out[k].size()), MPI_FLOAT, next, 100 + k, MPI_COMM_WORLD, &sr[k]);
  MPI_Wait(&sr[1], MPI_STATUS_IGNORE);
  MPI_Wait(&rr[2], MPI_STATUS_IGNORE);
  MPI_Waitall(3, sr, MPI_STATUSES_IGNORE);
  MPI_Waitall(3, rr, MPI_STATUSES_IGNORE);
  for (int k = 0; k < 3; ++k) REQUIRE(in[k].front() == float(prev * 10 + k) && in[k].back() == float(prev * 10 + k));
  // probe + get_count, then zero-length
  std::vector<int> payload(size_t(17 + rank), rank);
  MPI_Request s;
  MPI_Isend(payload.data(), int(payload.size()), MPI_INT, next, 7, MPI_COMM_WORLD, &s);
  MPI_Status st;
  MPI_Probe(prev, 7, MPI_COMM_WORLD, &st);
  int n = 0;
  MPI_Get_count(&st, MPI_INT, &n);
  REQUIRE(n == 17 + prev);
  std::vector<int> got(static_cast<size_t>(n), 0);
  MPI_Recv(got.data(), n, MPI_INT, prev, 7, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
  REQUIRE(got[0] == prev);
  MPI_Wait(&s, MPI_STATUS_IGNORE);
  int dummy = 0;
  MPI_Isend(&dummy, 0, MPI_INT, next, 9, MPI_COMM_WORLD, &s);
  MPI_Recv(&dummy, 0, MPI_INT, prev, 9, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
  MPI_Wait(&s, MPI_STATUS_IGNORE);
  // many messages: the descriptor ring wraps
  for (int i = 0; i < 3000; ++i) {
    int v = i, w = -1;
    MPI_Request a, b;
    MPI_Irecv(&w, 1, MPI_INT, prev, 1, MPI_COMM_WORLD, &b);
    MPI_Isend(&v, 1, MPI_INT, next, 1, MPI_COMM_WORLD, &a);
    MPI_Wait(&a, MPI_STATUS_IGNORE);
    MPI_Wait(&b, MPI_STATUS_IGNORE);
    REQUIRE(w == i);
  }
  MPI_Barrier(MPI_COMM_WORLD);
  if (rank == 0) std::printf("tz-mpi selftest ok (%d ranks)\n", size);
  MPI_Finalize();
  return 0;
}
