// comm.cu -- multi-GPU plumbing: one process per GPU, NCCL over NVLink/NVSwitch.
// (single-GPU build stage: communicator inactive)
#include "solver.cuh"

namespace rs {
bool comm_active() { return false; }
int comm_allreduce_sum(Ctx &, double *, int) { return RS_OK; }
int comm_allreduce_max(Ctx &, double *, int) { return RS_OK; }
}  // namespace rs

extern "C" int rs_comm_unique_id(char id[128]) { (void)id; rs::set_error("NCCL support not built"); return RS_ECOMM; }
extern "C" int rs_comm_init(int, int, const char *) { rs::set_error("NCCL support not built"); return RS_ECOMM; }
extern "C" int rs_comm_nranks(void) { return 1; }
extern "C" int rs_comm_rank(void) { return 0; }
extern "C" void rs_comm_finalize(void) {}
extern "C" void rs_slab_bounds(int d1, int nranks, int rank, int *j0, int *j1)
{
    const int base = d1 / nranks, rem = d1 % nranks;
    *j0 = rank * base + (rank < rem ? rank : rem);
    *j1 = *j0 + base + (rank < rem ? 1 : 0);
}
