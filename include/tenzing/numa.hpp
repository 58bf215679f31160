// NUMA placement of the calling thread and its pinned staging memory
// (reference: src/numa.cpp:12-109 — dead code there).
// No libnuma dependency: the node of a GPU is read from sysfs and the binding
// uses sched_setaffinity / the set_mempolicy syscall directly.
#pragma once

#include <string>
#include <vector>

namespace tenzing {
namespace numa {
// NUMA node of a PCI device ("0000:1b:00.0"), -1 if unknown
int node_of_pci_device(const std::string &pciBusId);
// CPUs of a node, empty if unknown
std::vector<int> cpus_of_node(int node);
// bind calling thread to the CPUs of `node` and prefer its memory; false if unsupported
bool bind_to_node(int node);
// bind to the node local to CUDA device `dev` (no-op + false without a GPU)
bool bind_to_local_memory(int cudaDevice);
} // namespace numa
} // namespace tenzing
