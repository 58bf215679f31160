// exhaustive search over distributed-SpMV schedules (reference: tenzing-dfs/examples/spmv.cu)
#include "spmv_run_strategy.hpp"
int main(int argc, char **argv) { return spmv_driver::main_impl(argc, argv, false, nullptr); }
