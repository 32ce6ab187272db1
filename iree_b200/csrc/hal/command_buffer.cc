// command_buffer.cc — recording and submission of dispatches, transfers and collectives.
//
// Stands in for runtime/src/iree/hal/drivers/cuda/{stream_command_buffer.c,
// graph_command_buffer.c}, hal/utils/{deferred_command_buffer.c, collective_batch.c} and
// the queue_execute/submit path of cuda_device.c:1168-1196,260-281.
//
// A command buffer is a recorded command list. In STREAM mode queue_execute replays it
// onto the dispatch stream (what the reference does with a deferred command buffer,
// hal/utils/deferred_work_queue.c:1198-1214 → stream_command_buffer.c); in GRAPH mode
// end() lowers the list to a CUgraph (kernel / memset / memcpy nodes, execution barriers
// as empty join nodes, graph_command_buffer.c:413-439) and queue_execute launches the
// instantiated graph.
//
// Kernel parameter packing is the reference's (stream_command_buffer.c:532-587): one
// device pointer per binding in layout order, then one slot per 32-bit push constant.
// For exports that ship a TMA recipe the encoded CUtensorMaps follow as trailing
// parameters (include/iree_b200_tma_recipe.h).

#include <algorithm>
#include <cstdlib>
#include <memory>

#include "../kernels/collective_config.h"
#include "internal.h"

using namespace b200hal;

namespace {

struct AlignedFree {
  void operator()(void* p) const { free(p); }
};

enum class CommandKind { kDispatch, kFill, kUpdate, kCopy, kCollective, kBarrier };

struct Command {
  CommandKind kind = CommandKind::kBarrier;
  // dispatch
  CUfunction function = nullptr;
  uint32_t grid[3] = {1, 1, 1};
  uint32_t block[3] = {1, 1, 1};
  uint32_t shared_memory = 0;
  std::unique_ptr<uint8_t, AlignedFree> storage;  // payload slots / captured bytes
  void** params = nullptr;                        // points into |storage|
  // fill / update / copy
  CUdeviceptr target = 0;
  CUdeviceptr source = 0;
  uint64_t length = 0;
  uint32_t pattern = 0;
  uint32_t pattern_length = 0;
  // collective
  iree_b200_channel_t* channel = nullptr;
  uint32_t op = 0;
  uint32_t param = 0;
  uint64_t element_count = 0;
  // Device memory the command reads / writes: [begin, end) byte ranges from its bindings
  // (hazard tracking of the dataflow replay, replay_dataflow).
  struct Range {
    uint64_t begin, end;
  };
  std::vector<Range> reads, writes;
  void touch(bool write, uint64_t ptr, uint64_t length) {
    if (ptr && length) (write ? writes : reads).push_back(Range{ptr, ptr + length});
  }
};

enum class State { kInitial, kRecording, kExecutable };

}  // namespace

struct iree_b200_command_buffer_t : b200hal::RefCounted {
  iree_b200_device_t* device = nullptr;
  int32_t mode = IREE_B200_COMMAND_BUFFER_MODE_STREAM;
  State state = State::kInitial;
  std::vector<Command> commands;
  // resource_set (stream_command_buffer.c:528-530,569-571)
  std::vector<iree_b200_buffer_t*> buffers;
  std::vector<iree_b200_executable_t*> executables;
  std::vector<iree_b200_channel_t*> channels;
  CUgraph graph = nullptr;
  CUgraphExec graph_exec = nullptr;
};

namespace {

iree_b200_status_t resolve_ref(iree_b200_command_buffer_t* cb, const iree_b200_buffer_ref_t& ref,
                               bool allow_null, CUdeviceptr* out_ptr, uint64_t* out_length) {
  *out_ptr = 0;
  *out_length = 0;
  if (!ref.buffer) {
    if (allow_null) return IREE_B200_STATUS_OK;
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "buffer reference is NULL");
  }
  const uint64_t total = ref.buffer->byte_length;
  if (ref.offset > total) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "binding offset %llu exceeds buffer of %llu bytes",
                       (unsigned long long)ref.offset, (unsigned long long)total);
  }
  uint64_t length = ref.length == UINT64_MAX ? total - ref.offset : ref.length;
  if (length > total - ref.offset) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "binding range [%llu, +%llu) exceeds buffer of %llu bytes",
                       (unsigned long long)ref.offset, (unsigned long long)length,
                       (unsigned long long)total);
  }
  *out_ptr = ref.buffer->device_ptr + ref.offset;
  *out_length = length;
  iree_b200_buffer_retain(ref.buffer);
  cb->buffers.push_back(ref.buffer);
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t require_recording(iree_b200_command_buffer_t* cb) {
  if (!cb) return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "command buffer is NULL");
  if (cb->state != State::kRecording) {
    return make_status(IREE_B200_STATUS_FAILED_PRECONDITION,
                       "command buffer is not in the recording state (begin/end mismatch)");
  }
  return IREE_B200_STATUS_OK;
}

uint32_t collective_element_size(uint32_t element_type) {
  switch (element_type) {
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_8:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_8:
      return 1;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_16:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_16:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_16:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_BFLOAT_16:
      return 2;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_32:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_32:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_32:
      return 4;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_64:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_64:
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_64:
      return 8;
    default:
      return 0;
  }
}

// nccl_channel.c:255-299 with BFLOAT_16 → ncclBfloat16.
iree_b200_status_t to_nccl_type(uint32_t element_type, ncclDataType_t* out) {
  switch (element_type) {
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_8: *out = ncclInt8; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_8: *out = ncclUint8; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_32: *out = ncclInt32; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_32: *out = ncclUint32; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_SINT_64: *out = ncclInt64; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_UINT_64: *out = ncclUint64; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_16: *out = ncclFloat16; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_32: *out = ncclFloat32; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_64: *out = ncclFloat64; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_BFLOAT_16: *out = ncclBfloat16; break;
    default:
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "unhandled element type for NCCL: %u", element_type);
  }
  return IREE_B200_STATUS_OK;
}

// nccl_channel.c:301-325.
iree_b200_status_t to_nccl_redop(uint32_t reduction, ncclRedOp_t* out) {
  switch (reduction) {
    case IREE_B200_COLLECTIVE_REDUCTION_SUM: *out = ncclSum; break;
    case IREE_B200_COLLECTIVE_REDUCTION_PRODUCT: *out = ncclProd; break;
    case IREE_B200_COLLECTIVE_REDUCTION_MINIMUM: *out = ncclMin; break;
    case IREE_B200_COLLECTIVE_REDUCTION_MAXIMUM: *out = ncclMax; break;
    case IREE_B200_COLLECTIVE_REDUCTION_AVERAGE: *out = ncclAvg; break;
    default:
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "unhandled reduction operation for NCCL: %u", reduction);
  }
  return IREE_B200_STATUS_OK;
}

// One entry of a batch (nccl_channel.c:327-540).
iree_b200_status_t issue_collective(const Command& c, CUstream stream) {
  const NcclSymbols* nccl = c.channel->nccl;
  const CudaSymbols* cu = c.channel->device->cu;
  ncclComm_t comm = c.channel->comm;
  const uint32_t kind = c.op & 0xff;
  const uint32_t reduction = (c.op >> 8) & 0xff;
  const uint32_t element_type = (c.op >> 16) & 0xff;
  ncclDataType_t dtype;
  B200_RETURN_IF_ERROR(to_nccl_type(element_type, &dtype));
  const void* send = reinterpret_cast<const void*>(c.source);
  void* recv = reinterpret_cast<void*>(c.target);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  switch (kind) {
    case IREE_B200_COLLECTIVE_KIND_ALL_GATHER:
      B200_NCCL(nccl, ncclAllGather(send, recv, c.element_count, dtype, comm, s));
      break;
    case IREE_B200_COLLECTIVE_KIND_ALL_REDUCE: {
      ncclRedOp_t redop;
      B200_RETURN_IF_ERROR(to_nccl_redop(reduction, &redop));
      B200_NCCL(nccl, ncclAllReduce(send, recv, c.element_count, dtype, redop, comm, s));
      break;
    }
    case IREE_B200_COLLECTIVE_KIND_ALL_TO_ALL: {
      // send/recv loop, element_count/count per peer (nccl_channel.c:381-410).
      const int count = c.channel->count;
      const uint64_t per_rank = c.element_count / (uint64_t)count;
      const uint64_t bytes = per_rank * collective_element_size(element_type);
      for (int r = 0; r < count; ++r) {
        B200_NCCL(nccl, ncclSend(static_cast<const uint8_t*>(send) + r * bytes, per_rank, dtype, r,
                                 comm, s));
        B200_NCCL(nccl,
                  ncclRecv(static_cast<uint8_t*>(recv) + r * bytes, per_rank, dtype, r, comm, s));
      }
      break;
    }
    case IREE_B200_COLLECTIVE_KIND_BROADCAST:
      B200_NCCL(nccl, ncclBroadcast(send, recv, c.element_count, dtype, (int)c.param, comm, s));
      break;
    case IREE_B200_COLLECTIVE_KIND_REDUCE: {
      ncclRedOp_t redop;
      B200_RETURN_IF_ERROR(to_nccl_redop(reduction, &redop));
      B200_NCCL(nccl, ncclReduce(send, recv, c.element_count, dtype, redop, (int)c.param, comm, s));
      break;
    }
    case IREE_B200_COLLECTIVE_KIND_REDUCE_SCATTER: {
      ncclRedOp_t redop;
      B200_RETURN_IF_ERROR(to_nccl_redop(reduction, &redop));
      B200_NCCL(nccl, ncclReduceScatter(send, recv, c.element_count, dtype, redop, comm, s));
      break;
    }
    case IREE_B200_COLLECTIVE_KIND_SEND:
      B200_NCCL(nccl, ncclSend(send, c.element_count, dtype, (int)c.param, comm, s));
      break;
    case IREE_B200_COLLECTIVE_KIND_RECV:
      B200_NCCL(nccl, ncclRecv(recv, c.element_count, dtype, (int)c.param, comm, s));
      break;
    case IREE_B200_COLLECTIVE_KIND_SEND_RECV: {
      // param: target rank in the low 16 bits, source rank in the high 16; -1 = none,
      // and a missing source zero-fills the receive buffer (nccl_channel.c:491-537).
      const int16_t send_rank = (int16_t)(c.param & 0xffff);
      const int16_t recv_rank = (int16_t)(c.param >> 16);
      if (send_rank != -1) {
        B200_NCCL(nccl, ncclSend(send, c.element_count, dtype, send_rank, comm, s));
      }
      if (recv_rank != -1) {
        B200_NCCL(nccl, ncclRecv(recv, c.element_count, dtype, recv_rank, comm, s));
      } else {
        B200_CU(cu, cuMemsetD8Async(c.target, 0,
                                    c.element_count * collective_element_size(element_type),
                                    stream));
      }
      break;
    }
    default:
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "unhandled collective kind %u", kind);
  }
  return IREE_B200_STATUS_OK;
}

// True when |c| can run on the symmetric-region kernels (kernels/collective_kernels.cu).
bool symmetric_all_reduce_eligible(const Command& c, int* out_type_index) {
  const auto& sym = c.channel->symmetric;
  if (!sym.attached || c.channel->count < 2) return false;
  const uint32_t kind = c.op & 0xff, reduction = (c.op >> 8) & 0xff, type = (c.op >> 16) & 0xff;
  if (kind != IREE_B200_COLLECTIVE_KIND_ALL_REDUCE || reduction != IREE_B200_COLLECTIVE_REDUCTION_SUM) {
    return false;
  }
  switch (type) {
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_16: *out_type_index = 0; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_BFLOAT_16: *out_type_index = 1; break;
    case IREE_B200_COLLECTIVE_ELEMENT_TYPE_FLOAT_32: *out_type_index = 2; break;
    default: return false;
  }
  const uint64_t bytes = c.element_count * collective_element_size(type);
  if (c.source != c.target || !bytes || bytes % 16 || c.target % 16) return false;
  return c.target >= sym.local_base && bytes <= sym.byte_length &&
         c.target - sym.local_base <= sym.byte_length - bytes;
}

// |chunks| > 1: one launch covers |chunks| equally sized ranges |chunk_stride| bytes apart;
// range c (c >= 1) is released by writing ready_base + c to the channel's ready word.
iree_b200_status_t launch_symmetric_all_reduce(const Command& c, int type_index, CUstream stream,
                                               uint32_t chunks = 1, uint64_t chunk_stride = 0,
                                               uint32_t ready_base = 0,
                                               CUDA_KERNEL_NODE_PARAMS* out_node_params = nullptr,
                                               CUgraph graph = nullptr,
                                               const CUgraphNode* deps = nullptr,
                                               size_t dep_count = 0, CUgraphNode* out_node = nullptr) {
  using Cfg = b200::CollectiveConfig;
  iree_b200_channel_t* channel = c.channel;
  const auto& sym = channel->symmetric;
  const CudaSymbols* cu = channel->device->cu;
  const uint32_t type = (c.op >> 16) & 0xff;
  uint64_t table = sym.table, multicast = sym.multicast_base;
  uint64_t byte_offset = c.target - sym.local_base;
  uint64_t vectors = c.element_count * collective_element_size(type) / 16;
  uint64_t ready = sym.table + Cfg::kReadyOffset;
  int rank = channel->rank, world = channel->count;
  uint32_t tuning = sym.tuning;
  const uint32_t threads = (uint32_t)sym.threads, unroll = std::max<uint32_t>(1, tuning & 0xff);
  // Every rank derives the same grid from the same element count (the flag words are per
  // block index, so the grids must agree across ranks).
  const uint64_t per_rank = (vectors + world - 1) / world;
  const uint64_t want = (per_rank + threads * unroll - 1) / (threads * unroll);
  const uint32_t blocks = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(want, sym.max_blocks));
  void* params[] = {&table, &multicast,  &byte_offset, &vectors, &chunks, &chunk_stride,
                    &ready, &ready_base, &rank,        &world,   &tuning};
  if (out_node_params) {
    // Graph command buffers: the same launch as a kernel node (cuGraphAddKernelNode copies
    // the parameter values, so stack storage is enough).
    memset(out_node_params, 0, sizeof(*out_node_params));
    out_node_params->func = sym.all_reduce_sum[type_index];
    out_node_params->gridDimX = blocks;
    out_node_params->gridDimY = out_node_params->gridDimZ = 1;
    out_node_params->blockDimX = threads;
    out_node_params->blockDimY = out_node_params->blockDimZ = 1;
    out_node_params->kernelParams = params;
    B200_CU(cu, cuGraphAddKernelNode(out_node, graph, deps, dep_count, out_node_params));
    return IREE_B200_STATUS_OK;
  }
  B200_CU(cu, cuLaunchKernel(sym.all_reduce_sum[type_index], blocks, 1, 1, threads, 1, 1, 0,
                             stream, params, nullptr));
  return IREE_B200_STATUS_OK;
}

// iree_hal_cuda_nccl_submit_batch (nccl_channel.c:542-588): one group per flushed batch.
// With a collective stream the batch is ordered after everything already on the dispatch
// stream and left running; |*pending_join| tells the caller to join before the next
// dependent command (the next execution barrier or the end of the command buffer).
iree_b200_status_t flush_collectives(iree_b200_device_t* device, std::vector<const Command*>* batch,
                                     bool* pending_join, CUevent fork_event) {
  if (batch->empty()) return IREE_B200_STATUS_OK;
  const CudaSymbols* cu = device->cu;
  CUstream stream = device->dispatch_stream;
  if (device->collective_stream) {
    B200_CU(cu, cuEventRecord(fork_event, device->dispatch_stream));
    B200_CU(cu, cuStreamWaitEvent(device->collective_stream, fork_event, 0));
    stream = device->collective_stream;
    *pending_join = true;
  }
  // In-place all_reduce(sum) inside the channel's symmetric region: our own NVLink kernels.
  std::vector<const Command*> for_nccl;
  for (const Command* c : *batch) {
    int type_index = -1;
    if (symmetric_all_reduce_eligible(*c, &type_index)) {
      iree_b200_status_t status = launch_symmetric_all_reduce(*c, type_index, stream);
      if (status != IREE_B200_STATUS_OK) {
        batch->clear();
        return status;
      }
      c->channel->symmetric_collectives.fetch_add(1);
    } else {
      for_nccl.push_back(c);
    }
  }
  batch->clear();
  if (for_nccl.empty()) return IREE_B200_STATUS_OK;
  const NcclSymbols* nccl = for_nccl.front()->channel->nccl;
  B200_NCCL(nccl, ncclGroupStart());
  iree_b200_status_t status = IREE_B200_STATUS_OK;
  for (const Command* c : for_nccl) {
    status = issue_collective(*c, stream);
    if (status != IREE_B200_STATUS_OK) break;
    c->channel->nccl_collectives.fetch_add(1);
  }
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  ncclResult_t end_result = nccl->ncclGroupEnd();
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  if (end_result != ncclSuccess) {
    return nccl_status(nccl, end_result, "ncclGroupEnd()", __FILE__, __LINE__);
  }
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t join_collectives(iree_b200_device_t* device, bool* pending_join,
                                    CUevent join_event) {
  if (!*pending_join) return IREE_B200_STATUS_OK;
  const CudaSymbols* cu = device->cu;
  B200_CU(cu, cuEventRecord(join_event, device->collective_stream));
  B200_CU(cu, cuStreamWaitEvent(device->dispatch_stream, join_event, 0));
  *pending_join = false;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t replay_dataflow(iree_b200_command_buffer_t* cb);

iree_b200_status_t replay_on_stream(iree_b200_command_buffer_t* cb) {
  iree_b200_device_t* device = cb->device;
  const CudaSymbols* cu = device->cu;
  if (device->collective_stream && device->params.overlap_collectives == 3) {
    return replay_dataflow(cb);
  }
  CUstream stream = device->dispatch_stream;
  std::vector<const Command*> batch;
  bool pending_join = false;
  CUevent fork_event = nullptr, join_event = nullptr;
  if (device->collective_stream) {
    B200_CU(cu, cuEventCreate(&fork_event, CU_EVENT_DISABLE_TIMING));
    B200_CU(cu, cuEventCreate(&join_event, CU_EVENT_DISABLE_TIMING));
  }
  // Range fusion (include/iree_b200_hal.h, overlap_collectives): chunk_of[i] = 0 for the
  // leader of a fused group (chunks_of[i] ranges, stride_of[i] bytes apart), c >= 1 for the
  // collective that releases range c of its group, -1 otherwise.
  const size_t n = cb->commands.size();
  std::vector<int32_t> chunk_of(n, -1), type_of(n, -1);
  std::vector<size_t> leader_of(n, 0);
  std::vector<uint32_t> base_of(n, 0);
  std::vector<uint32_t> chunks_of(n, 0);
  std::vector<uint64_t> stride_of(n, 0);
  if (device->collective_stream && device->params.overlap_collectives == 1) {
    auto single_eligible = [&](size_t i, int* type) {
      const Command& c = cb->commands[i];
      return c.kind == CommandKind::kCollective &&
             (i == 0 || cb->commands[i - 1].kind != CommandKind::kCollective) &&
             (i + 1 == n || cb->commands[i + 1].kind != CommandKind::kCollective) &&
             symmetric_all_reduce_eligible(c, type);
    };
    for (size_t i = 0; i < n; ++i) {
      int type = -1;
      if (chunk_of[i] != -1 || !single_eligible(i, &type)) continue;
      const Command& lead = cb->commands[i];
      std::vector<size_t> members{i};
      uint64_t stride = 0;
      for (size_t j = i + 1; j < n && members.size() < 64; ++j) {
        const Command& c = cb->commands[j];
        if (c.kind == CommandKind::kBarrier) break;
        if (c.kind != CommandKind::kCollective) continue;
        int type2 = -1;
        const Command& prev = cb->commands[members.back()];
        if (!single_eligible(j, &type2) || type2 != type || c.channel != lead.channel ||
            c.op != lead.op || c.element_count != lead.element_count || c.target <= prev.target) {
          break;
        }
        const uint64_t d = c.target - prev.target;
        if (members.size() == 1) stride = d;
        if (d != stride) break;
        members.push_back(j);
      }
      if (members.size() < 2) continue;
      for (size_t m = 0; m < members.size(); ++m) {
        chunk_of[members[m]] = (int32_t)m;
        leader_of[members[m]] = i;
      }
      chunks_of[i] = (uint32_t)members.size();
      stride_of[i] = stride;
      type_of[i] = type;
    }
  }
  iree_b200_status_t status = IREE_B200_STATUS_OK;
  // Fused launches whose ranges have not all been released yet: (leader index, released).
  std::vector<std::pair<size_t, uint32_t>> open_groups;
  auto run = [&]() -> iree_b200_status_t {
    for (size_t index = 0; index < n; ++index) {
      const Command& c = cb->commands[index];
      if (c.kind == CommandKind::kCollective && chunk_of[index] == 0) {
        // Leader: fork behind producer 0 and launch the persistent kernel for all ranges.
        auto& sym = c.channel->symmetric;
        const uint32_t ready_base = base_of[index] = sym.ready_sequence;
        sym.ready_sequence += chunks_of[index] - 1;
        B200_CU(cu, cuEventRecord(fork_event, device->dispatch_stream));
        B200_CU(cu, cuStreamWaitEvent(device->collective_stream, fork_event, 0));
        B200_RETURN_IF_ERROR(launch_symmetric_all_reduce(c, type_of[index],
                                                         device->collective_stream, chunks_of[index],
                                                         stride_of[index], ready_base));
        pending_join = true;
        open_groups.emplace_back(index, 0u);
        c.channel->symmetric_collectives.fetch_add(1);
        continue;
      }
      if (c.kind == CommandKind::kCollective && chunk_of[index] > 0) {
        // Range |chunk_of| is complete on this rank once everything queued so far is.
        const CUdeviceptr ready = c.channel->symmetric.table + b200::CollectiveConfig::kReadyOffset;
        B200_CU(cu, cuStreamWriteValue32(device->dispatch_stream, ready,
                                         base_of[leader_of[index]] + (uint32_t)chunk_of[index],
                                         CU_STREAM_WRITE_VALUE_DEFAULT));
        for (auto& g : open_groups) {
          if (g.first == leader_of[index]) g.second = (uint32_t)chunk_of[index];
        }
        c.channel->symmetric_collectives.fetch_add(1);
        continue;
      }
      if (c.kind == CommandKind::kCollective) {
        batch.push_back(&c);
        continue;
      }
      // Every other command flushes the batch first (stream_command_buffer.c:154-171).
      B200_RETURN_IF_ERROR(flush_collectives(device, &batch, &pending_join, fork_event));
      switch (c.kind) {
        case CommandKind::kBarrier:
          B200_RETURN_IF_ERROR(join_collectives(device, &pending_join, join_event));
          break;
        case CommandKind::kDispatch:
          B200_CU(cu, cuLaunchKernel(c.function, c.grid[0], c.grid[1], c.grid[2], c.block[0],
                                     c.block[1], c.block[2], c.shared_memory, stream, c.params,
                                     nullptr));
          break;
        case CommandKind::kFill:
          if (c.pattern_length == 4) {
            B200_CU(cu, cuMemsetD32Async(c.target, c.pattern, c.length / 4, stream));
          } else if (c.pattern_length == 2) {
            B200_CU(cu, cuMemsetD16Async(c.target, (uint16_t)c.pattern, c.length / 2, stream));
          } else {
            B200_CU(cu, cuMemsetD8Async(c.target, (uint8_t)c.pattern, c.length, stream));
          }
          break;
        case CommandKind::kUpdate:
          B200_CU(cu, cuMemcpyHtoDAsync(c.target, c.storage.get(), c.length, stream));
          break;
        case CommandKind::kCopy:
          B200_CU(cu, cuMemcpyAsync(c.target, c.source, c.length, stream));
          break;
        default:
          break;
      }
    }
    B200_RETURN_IF_ERROR(flush_collectives(device, &batch, &pending_join, fork_event));
    B200_RETURN_IF_ERROR(join_collectives(device, &pending_join, join_event));
    return IREE_B200_STATUS_OK;
  };
  status = run();
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  if (status != IREE_B200_STATUS_OK) {
    // A persistent all-reduce launched above may still be waiting for ranges whose release
    // was never queued: release them now (the data is garbage, the status says so) so that
    // the kernel retires instead of spinning under the caller's wait_idle.
    for (const auto& g : open_groups) {
      const Command& lead = cb->commands[g.first];
      const uint32_t last = chunks_of[g.first] - 1;
      if (g.second >= last) continue;
      cu->cuStreamWriteValue32(device->dispatch_stream,
                               lead.channel->symmetric.table + b200::CollectiveConfig::kReadyOffset,
                               base_of[g.first] + last, CU_STREAM_WRITE_VALUE_DEFAULT);
    }
  }
  if (fork_event) cu->cuEventDestroy(fork_event);
  if (join_event) cu->cuEventDestroy(join_event);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

// Launches one non-collective, non-barrier command on |stream|.
iree_b200_status_t launch_command(const CudaSymbols* cu, const Command& c, CUstream stream) {
  switch (c.kind) {
    case CommandKind::kDispatch:
      B200_CU(cu, cuLaunchKernel(c.function, c.grid[0], c.grid[1], c.grid[2], c.block[0], c.block[1],
                                 c.block[2], c.shared_memory, stream, c.params, nullptr));
      break;
    case CommandKind::kFill:
      if (c.pattern_length == 4) {
        B200_CU(cu, cuMemsetD32Async(c.target, c.pattern, c.length / 4, stream));
      } else if (c.pattern_length == 2) {
        B200_CU(cu, cuMemsetD16Async(c.target, (uint16_t)c.pattern, c.length / 2, stream));
      } else {
        B200_CU(cu, cuMemsetD8Async(c.target, (uint8_t)c.pattern, c.length, stream));
      }
      break;
    case CommandKind::kUpdate:
      B200_CU(cu, cuMemcpyHtoDAsync(c.target, c.storage.get(), c.length, stream));
      break;
    case CommandKind::kCopy:
      B200_CU(cu, cuMemcpyAsync(c.target, c.source, c.length, stream));
      break;
    default:
      break;
  }
  return IREE_B200_STATUS_OK;
}

bool ranges_overlap(const std::vector<Command::Range>& a, const std::vector<Command::Range>& b) {
  for (const auto& x : a) {
    for (const auto& y : b) {
      if (x.begin < y.end && y.begin < x.end) return true;
    }
  }
  return false;
}

// A hazard orders |later| behind |earlier|: read-after-write, write-after-read or
// write-after-write on overlapping device memory.
bool has_hazard(const Command& earlier, const Command& later) {
  return ranges_overlap(earlier.writes, later.reads) || ranges_overlap(earlier.writes, later.writes) ||
         ranges_overlap(earlier.reads, later.writes);
}

// overlap_collectives == 3: DATAFLOW replay on two in-order streams. Dispatches, fills and
// copies go to the dispatch stream and collectives to the collective stream, each in
// recording order; an execution barrier does not join the streams wholesale — a command
// waits (cuStreamWaitEvent) only for the latest command on the OTHER stream, recorded before
// the last barrier that precedes it, with which it has a data hazard on its binding ranges.
// Commands of one barrier-delimited group are unordered by contract; same-stream order is
// implicit. Kernels touch device memory through their bindings only (the dispatch ABI), so
// the recorded ranges are the complete access set and the result equals the fully
// serialised execution. The streams join at the end of the command buffer. This is what
// lets the all-reduce of chunk c of a row-parallel GEMM run under the GEMMs of other
// chunks while only its true consumer waits for it (iree_b200/tp_layer.py).
iree_b200_status_t replay_dataflow(iree_b200_command_buffer_t* cb) {
  iree_b200_device_t* device = cb->device;
  const CudaSymbols* cu = device->cu;
  const size_t n = cb->commands.size();
  CUstream streams[2] = {device->dispatch_stream, device->collective_stream};
  std::vector<int> stream_of(n, -1), group_of(n, 0);
  int group = 0;
  for (size_t i = 0; i < n; ++i) {
    const Command& c = cb->commands[i];
    if (c.kind == CommandKind::kBarrier) {
      ++group;
      continue;
    }
    group_of[i] = group;
    stream_of[i] = c.kind == CommandKind::kCollective ? 1 : 0;
  }
  // wait_for[j] = the command on the other stream j must wait for (or -1).
  std::vector<int64_t> wait_for(n, -1);
  std::vector<bool> signals(n, false);
  for (size_t j = 0; j < n; ++j) {
    if (stream_of[j] < 0) continue;
    for (size_t i = j; i-- > 0;) {
      if (stream_of[i] < 0 || stream_of[i] == stream_of[j] || group_of[i] == group_of[j]) continue;
      if (has_hazard(cb->commands[i], cb->commands[j])) {
        wait_for[j] = (int64_t)i;
        signals[i] = true;
        break;  // stream order covers every earlier command of that stream
      }
    }
  }
  std::vector<CUevent> events(n, nullptr);
  CUevent fork = nullptr, join = nullptr;
  int64_t last_waited[2] = {-1, -1};  // per stream: latest command of the other one already waited for
  auto run = [&]() -> iree_b200_status_t {
    // Everything already queued on the dispatch stream precedes the command buffer.
    B200_CU(cu, cuEventCreate(&fork, CU_EVENT_DISABLE_TIMING));
    B200_CU(cu, cuEventRecord(fork, streams[0]));
    B200_CU(cu, cuStreamWaitEvent(streams[1], fork, 0));
    bool used_collective_stream = false;
    for (size_t j = 0; j < n; ++j) {
      const Command& c = cb->commands[j];
      const int sj = stream_of[j];
      if (sj < 0) continue;
      if (wait_for[j] > last_waited[sj]) {
        B200_CU(cu, cuStreamWaitEvent(streams[sj], events[(size_t)wait_for[j]], 0));
        last_waited[sj] = wait_for[j];
      }
      if (c.kind == CommandKind::kCollective) {
        used_collective_stream = true;
        int type_index = -1;
        if (symmetric_all_reduce_eligible(c, &type_index)) {
          B200_RETURN_IF_ERROR(launch_symmetric_all_reduce(c, type_index, streams[1]));
          c.channel->symmetric_collectives.fetch_add(1);
        } else {
          const NcclSymbols* nccl = c.channel->nccl;
          B200_NCCL(nccl, ncclGroupStart());
          iree_b200_status_t status = issue_collective(c, streams[1]);
          std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
          ncclResult_t end_result = nccl->ncclGroupEnd();
          if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
          if (end_result != ncclSuccess) {
            return nccl_status(nccl, end_result, "ncclGroupEnd()", __FILE__, __LINE__);
          }
          c.channel->nccl_collectives.fetch_add(1);
        }
      } else {
        B200_RETURN_IF_ERROR(launch_command(cu, c, streams[0]));
      }
      if (signals[j]) {
        B200_CU(cu, cuEventCreate(&events[j], CU_EVENT_DISABLE_TIMING));
        B200_CU(cu, cuEventRecord(events[j], streams[sj]));
      }
    }
    if (used_collective_stream) {
      B200_CU(cu, cuEventCreate(&join, CU_EVENT_DISABLE_TIMING));
      B200_CU(cu, cuEventRecord(join, streams[1]));
      B200_CU(cu, cuStreamWaitEvent(streams[0], join, 0));
    }
    return IREE_B200_STATUS_OK;
  };
  iree_b200_status_t status = run();
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  // Destroying an event that is still pending only defers the release of its resources.
  for (CUevent e : events) {
    if (e) cu->cuEventDestroy(e);
  }
  if (fork) cu->cuEventDestroy(fork);
  if (join) cu->cuEventDestroy(join);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

// One batch of collectives (the commands between two barriers, as in stream mode) becomes
// one child-graph node: the NCCL calls are stream-captured on a scratch stream — the recipe
// the reference leaves as a TODO (graph_command_buffer.c:291-327) — and the captured graph
// is embedded with cuGraphAddChildGraphNode (which clones it).
iree_b200_status_t capture_collectives(iree_b200_command_buffer_t* cb,
                                       const std::vector<const Command*>& batch,
                                       const CUgraphNode* deps, size_t dep_count,
                                       CUgraphNode* out_node) {
  iree_b200_device_t* device = cb->device;
  const CudaSymbols* cu = device->cu;
  const NcclSymbols* nccl = batch.front()->channel->nccl;
  CUstream scratch = nullptr;
  B200_CU(cu, cuStreamCreate(&scratch, CU_STREAM_NON_BLOCKING));
  // RELAXED: NCCL may allocate or set up connections on a communicator's first collective.
  CUresult r = cu->cuStreamBeginCapture(scratch, CU_STREAM_CAPTURE_MODE_RELAXED);
  if (r != CUDA_SUCCESS) {
    cu->cuStreamDestroy(scratch);
    return cu_status(cu, r, "cuStreamBeginCapture", __FILE__, __LINE__);
  }
  iree_b200_status_t status = IREE_B200_STATUS_OK;
  std::string message;
  ncclResult_t group = nccl->ncclGroupStart();
  if (group != ncclSuccess) {
    status = nccl_status(nccl, group, "ncclGroupStart()", __FILE__, __LINE__);
  } else {
    for (const Command* c : batch) {
      status = issue_collective(*c, scratch);
      if (status != IREE_B200_STATUS_OK) break;
    }
    if (status != IREE_B200_STATUS_OK) message = iree_b200_status_message();
    group = nccl->ncclGroupEnd();
    if (status == IREE_B200_STATUS_OK && group != ncclSuccess) {
      status = nccl_status(nccl, group, "ncclGroupEnd()", __FILE__, __LINE__);
    }
  }
  if (status != IREE_B200_STATUS_OK && message.empty()) message = iree_b200_status_message();
  CUgraph child = nullptr;
  r = cu->cuStreamEndCapture(scratch, &child);  // always: it also ends a failed capture
  cu->cuStreamDestroy(scratch);
  if (status == IREE_B200_STATUS_OK && (r != CUDA_SUCCESS || !child)) {
    status = cu_status(cu, r, "cuStreamEndCapture", __FILE__, __LINE__);
    message = iree_b200_status_message();
  }
  if (status == IREE_B200_STATUS_OK) {
    r = cu->cuGraphAddChildGraphNode(out_node, cb->graph, deps, dep_count, child);
    if (r != CUDA_SUCCESS) {
      status = cu_status(cu, r, "cuGraphAddChildGraphNode", __FILE__, __LINE__);
      message = iree_b200_status_message();
    }
  }
  if (child) cu->cuGraphDestroy(child);
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  for (const Command* c : batch) c->channel->nccl_collectives.fetch_add(1);
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t build_graph(iree_b200_command_buffer_t* cb) {
  iree_b200_device_t* device = cb->device;
  const CudaSymbols* cu = device->cu;
  B200_CU(cu, cuGraphCreate(&cb->graph, 0));
  std::vector<CUgraphNode> group;   // nodes recorded since the last barrier
  CUgraphNode barrier = nullptr;    // join node of the previous barrier
  size_t node_count = 0;
  std::vector<const Command*> batch;  // collectives recorded since the last other command
  auto flush_batch = [&]() -> iree_b200_status_t {
    if (batch.empty()) return IREE_B200_STATUS_OK;
    // In-place all_reduce(sum) inside an attached symmetric region: a plain kernel node of
    // our NVLink kernel (where the reference has UNIMPLEMENTED, graph_command_buffer.c:
    // 291-327); everything else is captured from NCCL into one child graph.
    std::vector<const Command*> for_nccl;
    for (const Command* c : batch) {
      int type_index = -1;
      if (!symmetric_all_reduce_eligible(*c, &type_index)) {
        for_nccl.push_back(c);
        continue;
      }
      CUDA_KERNEL_NODE_PARAMS p;
      CUgraphNode node = nullptr;
      iree_b200_status_t status = launch_symmetric_all_reduce(
          *c, type_index, nullptr, 1, 0, 0, &p, cb->graph, barrier ? &barrier : nullptr,
          barrier ? 1 : 0, &node);
      if (status != IREE_B200_STATUS_OK) {
        batch.clear();
        return status;
      }
      c->channel->symmetric_collectives.fetch_add(1);
      group.push_back(node);
      ++node_count;
    }
    batch.clear();
    if (for_nccl.empty()) return IREE_B200_STATUS_OK;
    CUgraphNode node = nullptr;
    iree_b200_status_t status =
        capture_collectives(cb, for_nccl, barrier ? &barrier : nullptr, barrier ? 1 : 0, &node);
    if (status != IREE_B200_STATUS_OK) return status;
    group.push_back(node);
    ++node_count;
    return IREE_B200_STATUS_OK;
  };
  for (const Command& c : cb->commands) {
    if (c.kind == CommandKind::kCollective) {
      batch.push_back(&c);
      continue;
    }
    B200_RETURN_IF_ERROR(flush_batch());
    const CUgraphNode* deps = barrier ? &barrier : nullptr;
    const size_t dep_count = barrier ? 1 : 0;
    CUgraphNode node = nullptr;
    switch (c.kind) {
      case CommandKind::kBarrier: {
        if (group.empty()) break;
        B200_CU(cu, cuGraphAddEmptyNode(&node, cb->graph, group.data(), group.size()));
        barrier = node;
        group.clear();
        continue;
      }
      case CommandKind::kDispatch: {
        CUDA_KERNEL_NODE_PARAMS p;
        memset(&p, 0, sizeof(p));
        p.func = c.function;
        p.gridDimX = c.grid[0];
        p.gridDimY = c.grid[1];
        p.gridDimZ = c.grid[2];
        p.blockDimX = c.block[0];
        p.blockDimY = c.block[1];
        p.blockDimZ = c.block[2];
        p.sharedMemBytes = c.shared_memory;
        p.kernelParams = c.params;
        B200_CU(cu, cuGraphAddKernelNode(&node, cb->graph, deps, dep_count, &p));
        break;
      }
      case CommandKind::kFill: {
        CUDA_MEMSET_NODE_PARAMS p;
        memset(&p, 0, sizeof(p));
        p.dst = c.target;
        p.elementSize = c.pattern_length;
        p.value = c.pattern;
        p.width = c.length / c.pattern_length;
        p.height = 1;
        B200_CU(cu, cuGraphAddMemsetNode(&node, cb->graph, deps, dep_count, &p, device->cu_context));
        break;
      }
      case CommandKind::kUpdate:
      case CommandKind::kCopy: {
        CUDA_MEMCPY3D p;
        memset(&p, 0, sizeof(p));
        if (c.kind == CommandKind::kUpdate) {
          p.srcMemoryType = CU_MEMORYTYPE_HOST;
          p.srcHost = c.storage.get();
        } else {
          p.srcMemoryType = CU_MEMORYTYPE_DEVICE;
          p.srcDevice = c.source;
        }
        p.dstMemoryType = CU_MEMORYTYPE_DEVICE;
        p.dstDevice = c.target;
        p.WidthInBytes = c.length;
        p.Height = 1;
        p.Depth = 1;
        B200_CU(cu, cuGraphAddMemcpyNode(&node, cb->graph, deps, dep_count, &p, device->cu_context));
        break;
      }
      case CommandKind::kCollective:
        break;  // batched above
    }
    if (node) {
      group.push_back(node);
      ++node_count;
    }
  }
  B200_RETURN_IF_ERROR(flush_batch());
  // A command buffer that recorded nothing launchable stays without an executable graph
  // (queue_execute then only handles its semaphores).
  if (node_count) B200_CU(cu, cuGraphInstantiateWithFlags(&cb->graph_exec, cb->graph, 0));
  return IREE_B200_STATUS_OK;
}

}  // namespace

namespace b200hal {

void reap_in_flight(iree_b200_device_t* device, bool everything) {
  std::vector<iree_b200_device_t::InFlight> done;
  {
    std::lock_guard<std::mutex> lock(device->mutex);
    auto& list = device->in_flight;
    size_t kept = 0;
    for (size_t i = 0; i < list.size(); ++i) {
      // Submission order: once one is still running, so are the later ones.
      const bool finished =
          everything || (kept == 0 && device->cu->cuEventQuery(list[i].done) == CUDA_SUCCESS);
      if (finished) {
        done.push_back(list[i]);
      } else {
        list[kept++] = list[i];
      }
    }
    list.resize(kept);
  }
  // Outside the lock: releasing a command buffer may release the device.
  for (auto& e : done) {
    device->cu->cuEventDestroy(e.done);
    iree_b200_command_buffer_release(e.command_buffer);
  }
}

}  // namespace b200hal

extern "C" {

iree_b200_status_t iree_b200_command_buffer_create(iree_b200_device_t* device,
                                                   iree_b200_command_buffer_t** out_command_buffer) {
  B200_REQUIRE_ARG(device && out_command_buffer);
  auto* cb = new iree_b200_command_buffer_t();
  cb->device = device;
  cb->mode = device->params.command_buffer_mode;
  iree_b200_device_retain(device);
  *out_command_buffer = cb;
  return IREE_B200_STATUS_OK;
}

void iree_b200_command_buffer_retain(iree_b200_command_buffer_t* cb) {
  if (cb) cb->ref_count.fetch_add(1);
}

void iree_b200_command_buffer_release(iree_b200_command_buffer_t* cb) {
  if (!cb || cb->ref_count.fetch_sub(1) != 1) return;
  iree_b200_device_t* device = cb->device;
  device->cu->cuCtxSetCurrent(device->cu_context);
  if (cb->graph_exec) device->cu->cuGraphExecDestroy(cb->graph_exec);
  if (cb->graph) device->cu->cuGraphDestroy(cb->graph);
  for (auto* b : cb->buffers) iree_b200_buffer_release(b);
  for (auto* e : cb->executables) iree_b200_executable_release(e);
  for (auto* c : cb->channels) iree_b200_channel_release(c);
  delete cb;
  iree_b200_device_release(device);
}

iree_b200_status_t iree_b200_command_buffer_begin(iree_b200_command_buffer_t* cb) {
  B200_REQUIRE_ARG(cb);
  if (cb->state != State::kInitial) {
    return make_status(IREE_B200_STATUS_FAILED_PRECONDITION,
                       "command buffer is already recording or has been recorded");
  }
  cb->state = State::kRecording;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_command_buffer_end(iree_b200_command_buffer_t* cb) {
  B200_RETURN_IF_ERROR(require_recording(cb));
  if (cb->mode == IREE_B200_COMMAND_BUFFER_MODE_GRAPH) {
    cb->device->cu->cuCtxSetCurrent(cb->device->cu_context);
    B200_RETURN_IF_ERROR(build_graph(cb));
  }
  cb->state = State::kExecutable;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_command_buffer_execution_barrier(iree_b200_command_buffer_t* cb) {
  B200_RETURN_IF_ERROR(require_recording(cb));
  Command c;
  c.kind = CommandKind::kBarrier;
  cb->commands.push_back(std::move(c));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_command_buffer_fill_buffer(iree_b200_command_buffer_t* cb,
                                                        iree_b200_buffer_ref_t target,
                                                        const void* pattern,
                                                        size_t pattern_length) {
  B200_RETURN_IF_ERROR(require_recording(cb));
  B200_REQUIRE_ARG(pattern);
  if (pattern_length != 1 && pattern_length != 2 && pattern_length != 4) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "fill patterns must be 1, 2 or 4 bytes (got %zu)", pattern_length);
  }
  Command c;
  c.kind = CommandKind::kFill;
  B200_RETURN_IF_ERROR(resolve_ref(cb, target, false, &c.target, &c.length));
  if (c.target % pattern_length != 0 || c.length % pattern_length != 0) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "fill range must be aligned to the pattern length %zu", pattern_length);
  }
  memcpy(&c.pattern, pattern, pattern_length);
  c.pattern_length = (uint32_t)pattern_length;
  c.touch(true, c.target, c.length);
  if (c.length) cb->commands.push_back(std::move(c));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_command_buffer_update_buffer(iree_b200_command_buffer_t* cb,
                                                          const void* source, size_t source_offset,
                                                          iree_b200_buffer_ref_t target) {
  B200_RETURN_IF_ERROR(require_recording(cb));
  B200_REQUIRE_ARG(source);
  Command c;
  c.kind = CommandKind::kUpdate;
  B200_RETURN_IF_ERROR(resolve_ref(cb, target, false, &c.target, &c.length));
  if (!c.length) return IREE_B200_STATUS_OK;
  // The source is captured: callers may reuse their memory as soon as this returns
  // (stream_command_buffer.c:383-392).
  c.storage.reset(static_cast<uint8_t*>(aligned_alloc(64, (c.length + 63) / 64 * 64)));
  if (!c.storage) return make_status(IREE_B200_STATUS_RESOURCE_EXHAUSTED, "out of host memory");
  memcpy(c.storage.get(), static_cast<const uint8_t*>(source) + source_offset, c.length);
  c.touch(true, c.target, c.length);
  cb->commands.push_back(std::move(c));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_command_buffer_copy_buffer(iree_b200_command_buffer_t* cb,
                                                        iree_b200_buffer_ref_t source,
                                                        iree_b200_buffer_ref_t target) {
  B200_RETURN_IF_ERROR(require_recording(cb));
  Command c;
  c.kind = CommandKind::kCopy;
  uint64_t source_length = 0;
  B200_RETURN_IF_ERROR(resolve_ref(cb, source, false, &c.source, &source_length));
  B200_RETURN_IF_ERROR(resolve_ref(cb, target, false, &c.target, &c.length));
  if (source_length != c.length) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "copy source (%llu bytes) and target (%llu bytes) lengths differ",
                       (unsigned long long)source_length, (unsigned long long)c.length);
  }
  c.touch(false, c.source, c.length);
  c.touch(true, c.target, c.length);
  if (c.length) cb->commands.push_back(std::move(c));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_command_buffer_dispatch(
    iree_b200_command_buffer_t* cb, iree_b200_executable_t* executable, int32_t export_ordinal,
    const uint32_t workgroup_count[3], const uint32_t* constants, size_t constant_count,
    const iree_b200_buffer_ref_t* bindings, size_t binding_count, uint64_t flags) {
  B200_RETURN_IF_ERROR(require_recording(cb));
  B200_REQUIRE_ARG(executable && workgroup_count);
  B200_REQUIRE_ARG(constants || !constant_count);
  B200_REQUIRE_ARG(bindings || !binding_count);
  if (flags != 0) {
    // Indirect / custom-argument dispatches are rejected (stream_command_buffer.c:495-503).
    return make_status(IREE_B200_STATUS_UNIMPLEMENTED, "dispatch flags 0x%llx are not supported",
                       (unsigned long long)flags);
  }
  if (export_ordinal < 0 || (size_t)export_ordinal >= executable->exports.size()) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "export ordinal %d out of range; executable has %zu exports",
                       export_ordinal, executable->exports.size());
  }
  const KernelParams& k = executable->exports[export_ordinal];
  if (constant_count != k.constant_count) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "dispatch of `%s` requires %u constants but %zu were provided", k.name.c_str(),
                       k.constant_count, constant_count);
  }
  if (binding_count != k.binding_count) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "dispatch of `%s` requires %u bindings but %zu were provided", k.name.c_str(),
                       k.binding_count, binding_count);
  }
  // A zero-sized grid is a no-op (hal/command_buffer.c treats it as such).
  if (!workgroup_count[0] || !workgroup_count[1] || !workgroup_count[2]) {
    return IREE_B200_STATUS_OK;
  }

  const uint32_t map_count = k.has_recipe ? k.recipe.map_count : 0;
  const size_t slot_count = binding_count + constant_count;
  const size_t maps_bytes = (size_t)map_count * sizeof(CUtensorMap);
  const size_t slots_bytes = slot_count * 8;
  const size_t params_bytes = (slot_count + map_count) * sizeof(void*);
  const size_t total = (maps_bytes + slots_bytes + params_bytes + 63) / 64 * 64 + 64;

  Command c;
  c.kind = CommandKind::kDispatch;
  c.function = k.function;
  for (int d = 0; d < 3; ++d) {
    c.grid[d] = workgroup_count[d];
    c.block[d] = k.block_dims[d];
  }
  c.shared_memory = k.block_shared_memory_size;
  c.storage.reset(static_cast<uint8_t*>(aligned_alloc(64, total)));
  if (!c.storage) return make_status(IREE_B200_STATUS_RESOURCE_EXHAUSTED, "out of host memory");
  memset(c.storage.get(), 0, total);
  CUtensorMap* maps = reinterpret_cast<CUtensorMap*>(c.storage.get());
  uint64_t* slots = reinterpret_cast<uint64_t*>(c.storage.get() + maps_bytes);
  c.params = reinterpret_cast<void**>(c.storage.get() + maps_bytes + slots_bytes);

  uint64_t binding_ptrs[IREE_B200_MAX_DISPATCH_BINDING_COUNT] = {0};
  for (size_t i = 0; i < binding_count; ++i) {
    CUdeviceptr ptr = 0;
    uint64_t length = 0;
    B200_RETURN_IF_ERROR(resolve_ref(cb, bindings[i], /*allow_null=*/true, &ptr, &length));
    binding_ptrs[i] = ptr;
    slots[i] = ptr;
    c.params[i] = &slots[i];
    // cuda_executable_def.fbs BindingBits: 1 = READ_ONLY; anything else may be written.
    c.touch((k.binding_flags[i] & 1) == 0, ptr, length);
  }
  for (size_t i = 0; i < constant_count; ++i) {
    slots[binding_count + i] = constants[i];  // low 32 bits are the value
    c.params[binding_count + i] = &slots[binding_count + i];
  }
  if (map_count) {
    cb->device->cu->cuCtxSetCurrent(cb->device->cu_context);
    for (uint32_t m = 0; m < map_count; ++m) {
      const iree_b200_tma_map_t& map = k.recipe.maps[m];
      B200_RETURN_IF_ERROR(encode_tensor_map(cb->device->cu, map, binding_ptrs[map.binding],
                                             constants, constant_count, &maps[m]));
      c.params[slot_count + m] = &maps[m];
    }
  }
  iree_b200_executable_retain(executable);
  cb->executables.push_back(executable);
  cb->commands.push_back(std::move(c));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_command_buffer_collective(
    iree_b200_command_buffer_t* cb, iree_b200_channel_t* channel, uint32_t packed_op,
    uint32_t param, iree_b200_buffer_ref_t send_binding, iree_b200_buffer_ref_t recv_binding,
    uint64_t element_count) {
  B200_RETURN_IF_ERROR(require_recording(cb));
  B200_REQUIRE_ARG(channel);
  const uint32_t kind = packed_op & 0xff;
  const uint32_t element_type = (packed_op >> 16) & 0xff;
  if (kind > IREE_B200_COLLECTIVE_KIND_SEND_RECV) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "unhandled collective kind %u", kind);
  }
  const uint32_t element_size = collective_element_size(element_type);
  if (!element_size) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "unhandled collective element type %u",
                       element_type);
  }
  ncclDataType_t dtype;
  B200_RETURN_IF_ERROR(to_nccl_type(element_type, &dtype));
  Command c;
  c.kind = CommandKind::kCollective;
  c.channel = channel;
  c.op = packed_op;
  c.param = param;
  c.element_count = element_count;
  uint64_t send_length = 0, recv_length = 0;
  B200_RETURN_IF_ERROR(resolve_ref(cb, send_binding, true, &c.source, &send_length));
  B200_RETURN_IF_ERROR(resolve_ref(cb, recv_binding, true, &c.target, &recv_length));
  // Binding requirements per kind (command_buffer.h:300-400).
  const bool needs_send = kind != IREE_B200_COLLECTIVE_KIND_RECV &&
                          kind != IREE_B200_COLLECTIVE_KIND_SEND_RECV;
  const bool needs_recv = kind != IREE_B200_COLLECTIVE_KIND_SEND &&
                          kind != IREE_B200_COLLECTIVE_KIND_SEND_RECV;
  if ((needs_send && !c.source) || (needs_recv && !c.target)) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                       "collective kind %u is missing a required binding", kind);
  }
  uint64_t send_elements = element_count, recv_elements = element_count;
  if (kind == IREE_B200_COLLECTIVE_KIND_ALL_GATHER) recv_elements *= (uint64_t)channel->count;
  if (kind == IREE_B200_COLLECTIVE_KIND_REDUCE_SCATTER) send_elements *= (uint64_t)channel->count;
  if (c.source && send_length < send_elements * element_size) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "collective send binding holds %llu bytes but %llu are required",
                       (unsigned long long)send_length,
                       (unsigned long long)(send_elements * element_size));
  }
  if (c.target && recv_length < recv_elements * element_size) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "collective recv binding holds %llu bytes but %llu are required",
                       (unsigned long long)recv_length,
                       (unsigned long long)(recv_elements * element_size));
  }
  c.touch(false, c.source, send_elements * element_size);
  c.touch(true, c.target, recv_elements * element_size);
  iree_b200_channel_retain(channel);
  cb->channels.push_back(channel);
  cb->commands.push_back(std::move(c));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_device_queue_execute(
    iree_b200_device_t* device, size_t wait_count, iree_b200_semaphore_t* const* wait_semaphores,
    const uint64_t* wait_values, size_t signal_count,
    iree_b200_semaphore_t* const* signal_semaphores, const uint64_t* signal_values,
    iree_b200_command_buffer_t* command_buffer) {
  B200_REQUIRE_ARG(device);
  B200_REQUIRE_ARG(!wait_count || (wait_semaphores && wait_values));
  B200_REQUIRE_ARG(!signal_count || (signal_semaphores && signal_values));
  if (command_buffer && command_buffer->state != State::kExecutable) {
    return make_status(IREE_B200_STATUS_FAILED_PRECONDITION,
                       "command buffer has not been ended (state is not executable)");
  }
  const CudaSymbols* cu = device->cu;
  b200hal::reap_in_flight(device, /*everything=*/false);
  std::lock_guard<std::mutex> lock(device->mutex);
  B200_CU(cu, cuCtxSetCurrent(device->cu_context));

  // An operation that cannot be queued fails its signal semaphores, so that waiters
  // downstream see the failure instead of hanging (cuda_device.c:1187-1191).
  auto fail_signals = [&](iree_b200_status_t code, const std::string& message) {
    for (size_t i = 0; i < signal_count; ++i) {
      iree_b200_semaphore_t* s = signal_semaphores[i];
      std::lock_guard<std::mutex> sem_lock(s->mutex);
      if (s->failure == IREE_B200_STATUS_OK) {
        s->failure = code;
        s->failure_message = message;
      }
    }
  };

  // Device-side waits (cuda_device.c:148-158,229-240).
  for (size_t i = 0; i < wait_count; ++i) {
    iree_b200_semaphore_t* s = wait_semaphores[i];
    std::unique_lock<std::mutex> sem_lock(s->mutex);
    if (s->failure != IREE_B200_STATUS_OK) {
      const iree_b200_status_t code = s->failure;
      const std::string message = "wait semaphore has failed: " + s->failure_message;
      sem_lock.unlock();
      fail_signals(code, message);
      return make_status(code, "%s", message.c_str());
    }
    if (s->host_value >= wait_values[i]) continue;
    CUevent event = nullptr;
    for (auto& tp : s->timepoints) {
      if (tp.value >= wait_values[i]) {
        event = tp.event;
        break;
      }
    }
    if (!event) {
      // No deferred work queue here (the reference's supports wait-before-signal,
      // hal/utils/deferred_work_queue.c): submissions must be queued in dependency order.
      char message[128];
      snprintf(message, sizeof(message),
               "wait on semaphore value %llu that no queued operation signals",
               (unsigned long long)wait_values[i]);
      sem_lock.unlock();
      fail_signals(IREE_B200_STATUS_FAILED_PRECONDITION, message);
      return make_status(IREE_B200_STATUS_FAILED_PRECONDITION, "%s", message);
    }
    B200_CU(cu, cuStreamWaitEvent(device->dispatch_stream, event, 0));
  }

  iree_b200_status_t status = IREE_B200_STATUS_OK;
  if (command_buffer) {
    if (command_buffer->mode == IREE_B200_COMMAND_BUFFER_MODE_GRAPH) {
      if (command_buffer->graph_exec) {
        CUresult r = cu->cuGraphLaunch(command_buffer->graph_exec, device->dispatch_stream);
        if (r != CUDA_SUCCESS) status = cu_status(cu, r, "cuGraphLaunch", __FILE__, __LINE__);
      }
    } else {
      status = replay_on_stream(command_buffer);
    }
  }
  std::string message = status != IREE_B200_STATUS_OK ? iree_b200_status_message() : "";
  if (command_buffer && status == IREE_B200_STATUS_OK) {
    // Keep the command buffer (and through it its buffers, executables, captured update
    // bytes and graph) alive until the GPU is done with this submission.
    CUevent done = nullptr;
    if (cu->cuEventCreate(&done, CU_EVENT_DISABLE_TIMING) == CUDA_SUCCESS) {
      if (cu->cuEventRecord(done, device->dispatch_stream) == CUDA_SUCCESS) {
        iree_b200_command_buffer_retain(command_buffer);
        device->in_flight.push_back({command_buffer, done});
      } else {
        cu->cuEventDestroy(done);
      }
    }
  }

  for (size_t i = 0; i < signal_count; ++i) {
    iree_b200_semaphore_t* s = signal_semaphores[i];
    std::lock_guard<std::mutex> sem_lock(s->mutex);
    if (status != IREE_B200_STATUS_OK) {
      // Async failure delivery (cuda_device.c:1187-1191).
      s->failure = status;
      s->failure_message = message;
      continue;
    }
    CUevent event = nullptr;
    CUresult r = cu->cuEventCreate(&event, CU_EVENT_DISABLE_TIMING);
    if (r == CUDA_SUCCESS) r = cu->cuEventRecord(event, device->dispatch_stream);
    if (r != CUDA_SUCCESS) {
      status = cu_status(cu, r, "cuEventCreate/Record", __FILE__, __LINE__);
      message = iree_b200_status_message();
      s->failure = status;
      s->failure_message = message;
      continue;
    }
    s->timepoints.push_back({signal_values[i], event});
  }
  if (status != IREE_B200_STATUS_OK) return make_status(status, "%s", message.c_str());
  return IREE_B200_STATUS_OK;
}

}  // extern "C"
