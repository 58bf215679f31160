// MCTS (Coverage) over halo-exchange schedules (reference: tenzing-mcts/examples/halo_*.cu)
#include "halo_run_strategy.hpp"
int main(int argc, char **argv) { return halo_driver::main_impl(argc, argv, "Coverage"); }
