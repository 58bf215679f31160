// MCTS (Random) over distributed-SpMV schedules (reference: tenzing-mcts/examples/spmv_*.cu)
#include "spmv_run_strategy.hpp"
int main(int argc, char **argv) { return spmv_driver::main_impl(argc, argv, true, "Random"); }
