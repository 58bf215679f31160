#include "tenzing/numa.hpp"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include <sched.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <cuda_runtime_api.h>

namespace tenzing {
namespace numa {

int node_of_pci_device(const std::string &pciBusId) {
  std::string id = pciBusId;
  for (auto &c : id) c = static_cast<char>(std::tolower(c));
  std::ifstream f("/sys/bus/pci/devices/" + id + "/numa_node");
  int node = -1;
  if (f) f >> node;
  return node;
}

// parse "0-15,32-47"
std::vector<int> cpus_of_node(int node) {
  std::vector<int> cpus;
  if (node < 0) return cpus;
  std::ifstream f("/sys/devices/system/node/node" + std::to_string(node) + "/cpulist");
  std::string line;
  if (!f || !std::getline(f, line)) return cpus;
  std::stringstream ss(line);
  std::string part;
  while (std::getline(ss, part, ',')) {
    int a = 0, b = 0;
    if (std::sscanf(part.c_str(), "%d-%d", &a, &b) == 2) {
      for (int c = a; c <= b; ++c) cpus.push_back(c);
    } else if (std::sscanf(part.c_str(), "%d", &a) == 1) {
      cpus.push_back(a);
    }
  }
  return cpus;
}

bool bind_to_node(int node) {
  const std::vector<int> cpus = cpus_of_node(node);
  if (cpus.empty()) return false;
  cpu_set_t set;
  CPU_ZERO(&set);
  for (int c : cpus)
    if (c < CPU_SETSIZE) CPU_SET(c, &set);
  if (sched_setaffinity(0, sizeof(set), &set) != 0) return false;
#ifdef SYS_set_mempolicy
  // MPOL_PREFERRED = 1: allocate on `node` when possible (pinned staging buffers
  // created after this call land next to the GPU)
  unsigned long mask[16] = {0};
  if (node < int(sizeof(mask) * 8)) {
    mask[node / (8 * sizeof(unsigned long))] |= 1UL << (node % (8 * sizeof(unsigned long)));
    (void)syscall(SYS_set_mempolicy, 1, mask, sizeof(mask) * 8);
  }
#endif
  return true;
}

bool bind_to_local_memory(int cudaDevice) {
  char busId[32] = {0};
  if (cudaDeviceGetPCIBusId(busId, sizeof(busId), cudaDevice) != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return bind_to_node(node_of_pci_device(busId));
}

} // namespace numa
} // namespace tenzing
