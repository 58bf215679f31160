// Umbrella header + the namespace spellings used by the reference's docs
// (docs/api.md:16-25 writes SDP::BaseOp, SDP::Graph, ...; the code itself
// keeps most names at global scope).
#pragma once

#include "tenzing/benchmarker.hpp"
#include "tenzing/cast.hpp"
#include "tenzing/communicator.hpp"
#include "tenzing/counters.hpp"
#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/decision.hpp"
#include "tenzing/dfs.hpp"
#include "tenzing/event_synchronizer.hpp"
#include "tenzing/flag_synchronizer.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/init.hpp"
#include "tenzing/mcts.hpp"
#include "tenzing/numa.hpp"
#include "tenzing/numeric.hpp"
#include "tenzing/operation.hpp"
#include "tenzing/operation_compound.hpp"
#include "tenzing/operation_serdes.hpp"
#include "tenzing/host_exec.hpp"
#include "tenzing/host_streams.hpp"
#include "tenzing/ops_comm.hpp"
#include "tenzing/platform.hpp"
#include "tenzing/randomness.hpp"
#include "tenzing/reproduce.hpp"
#include "tenzing/schedule.hpp"
#include "tenzing/sequence.hpp"
#include "tenzing/state.hpp"
#include "tenzing/trap.hpp"

namespace tenzing {
namespace SDP {
using BaseOp = ::tenzing::OpBase;
using OpBase = ::tenzing::OpBase;
using BoundOp = ::tenzing::BoundOp;
using CpuOp = ::tenzing::CpuOp;
using GpuOp = ::tenzing::GpuOp;
using BoundGpuOp = ::tenzing::BoundGpuOp;
using CompoundOp = ::tenzing::CompoundOp;
using ChoiceOp = ::tenzing::ChoiceOp;
using Graph = ::tenzing::Graph<::tenzing::OpBase>;
} // namespace SDP
} // namespace tenzing
