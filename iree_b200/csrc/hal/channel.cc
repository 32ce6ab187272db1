// channel.cc — collective channels over NCCL and the MPI-free default channel provider.
//
// Stands in for runtime/src/iree/hal/drivers/cuda/nccl_channel.c:76-231, the channel
// creation path of cuda_device.c:845-927 and hal/utils/mpi_channel_provider.c. One rank
// per process per device, one communicator per channel, exactly as in the reference
// ("exactly one participant is allowed in a channel", cuda_device.c:858-868). Neither
// machine has MPI, so the default provider takes rank/count from the launcher's
// environment and exchanges the 128-byte id through a file.

#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdlib>
#include <thread>

#include "../kernels/collective_config.h"
#include "internal.h"

using namespace b200hal;

static_assert(sizeof(ncclUniqueId) == IREE_B200_CHANNEL_ID_SIZE, "id size");

namespace {

bool env_int(const char* name, int32_t* out) {
  const char* v = getenv(name);
  if (!v || !*v) return false;
  char* end = nullptr;
  long parsed = strtol(v, &end, 10);
  if (end == v || *end) return false;
  *out = (int32_t)parsed;
  return true;
}

}  // namespace

extern "C" {

iree_b200_status_t iree_b200_channel_id_create(uint8_t out_id[IREE_B200_CHANNEL_ID_SIZE]) {
  B200_REQUIRE_ARG(out_id);
  const NcclSymbols* nccl = nullptr;
  B200_RETURN_IF_ERROR(nccl_symbols(&nccl));
  ncclUniqueId id;
  memset(&id, 0, sizeof(id));
  B200_NCCL(nccl, ncclGetUniqueId(&id));
  memcpy(out_id, &id, sizeof(id));
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_channel_provider_query_default_rank_and_count(int32_t* out_rank,
                                                                           int32_t* out_count) {
  B200_REQUIRE_ARG(out_rank && out_count);
  static const char* const kPairs[][2] = {
      {"IREE_B200_RANK", "IREE_B200_WORLD_SIZE"},
      {"RANK", "WORLD_SIZE"},
      {"OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE"},
      {"PMI_RANK", "PMI_SIZE"},
  };
  for (const auto& pair : kPairs) {
    int32_t rank = 0, count = 0;
    if (env_int(pair[0], &rank) && env_int(pair[1], &count)) {
      if (count < 1 || rank < 0 || rank >= count) {
        return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "%s=%d / %s=%d is not a valid rank",
                           pair[0], rank, pair[1], count);
      }
      *out_rank = rank;
      *out_count = count;
      return IREE_B200_STATUS_OK;
    }
  }
  // No launcher: a single-participant group (mpi_channel_provider.c falls back the same
  // way when MPI is not initialised).
  *out_rank = 0;
  *out_count = 1;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_channel_provider_exchange_default_id(
    int32_t rank, uint8_t inout_id[IREE_B200_CHANNEL_ID_SIZE]) {
  B200_REQUIRE_ARG(inout_id);
  const char* path = getenv("IREE_B200_CHANNEL_ID_FILE");
  if (!path || !*path) {
    return make_status(IREE_B200_STATUS_FAILED_PRECONDITION,
                       "no channel id given and IREE_B200_CHANNEL_ID_FILE is not set; cannot "
                       "exchange the default id between ranks");
  }
  if (rank == 0) {
    // Write-then-rename so readers never observe a partial id.
    std::string tmp = std::string(path) + ".tmp." + std::to_string((long)getpid());
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return make_status(IREE_B200_STATUS_PERMISSION_DENIED, "cannot write %s", tmp.c_str());
    const size_t written = fwrite(inout_id, 1, IREE_B200_CHANNEL_ID_SIZE, f);
    fclose(f);
    if (written != IREE_B200_CHANNEL_ID_SIZE || rename(tmp.c_str(), path) != 0) {
      return make_status(IREE_B200_STATUS_DATA_LOSS, "cannot publish channel id at %s", path);
    }
    return IREE_B200_STATUS_OK;
  }
  int32_t timeout_s = 120;
  env_int("IREE_B200_CHANNEL_ID_TIMEOUT_S", &timeout_s);
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(timeout_s);
  for (;;) {
    struct stat st;
    if (stat(path, &st) == 0 && st.st_size == IREE_B200_CHANNEL_ID_SIZE) {
      FILE* f = fopen(path, "rb");
      if (f) {
        const size_t got = fread(inout_id, 1, IREE_B200_CHANNEL_ID_SIZE, f);
        fclose(f);
        if (got == IREE_B200_CHANNEL_ID_SIZE) return IREE_B200_STATUS_OK;
      }
    }
    if (std::chrono::steady_clock::now() >= deadline) {
      return make_status(IREE_B200_STATUS_DEADLINE_EXCEEDED,
                         "rank %d timed out waiting for the channel id at %s", rank, path);
    }
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
  }
}

iree_b200_status_t iree_b200_channel_create(iree_b200_device_t* device,
                                            const iree_b200_channel_params_t* params,
                                            iree_b200_channel_t** out_channel) {
  B200_REQUIRE_ARG(device && params && out_channel);
  *out_channel = nullptr;
  const NcclSymbols* nccl = nullptr;
  // cuda_device.c:849-856: NCCL must be loadable.
  B200_RETURN_IF_ERROR(nccl_symbols(&nccl));

  // Rank / count defaults from the provider (cuda_device.c:872-879).
  int32_t rank = params->rank, count = params->count;
  if (rank == IREE_B200_CHANNEL_RANK_DEFAULT || count == IREE_B200_CHANNEL_COUNT_DEFAULT) {
    int32_t default_rank = 0, default_count = 1;
    B200_RETURN_IF_ERROR(
        iree_b200_channel_provider_query_default_rank_and_count(&default_rank, &default_count));
    if (rank == IREE_B200_CHANNEL_RANK_DEFAULT) rank = default_rank;
    if (count == IREE_B200_CHANNEL_COUNT_DEFAULT) count = default_count;
  }
  if (count < 1 || rank < 0 || rank >= count) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "channel rank %d of %d is invalid", rank,
                       count);
  }

  // Id: given, or created on rank 0 and exchanged (cuda_device.c:885-903).
  ncclUniqueId id;
  memset(&id, 0, sizeof(id));
  if (params->id && params->id_length) {
    if (params->id_length != IREE_B200_CHANNEL_ID_SIZE) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "NCCL ID must be %d bytes matching the ncclUniqueId struct but caller "
                         "provided %zu bytes",
                         IREE_B200_CHANNEL_ID_SIZE, params->id_length);
    }
    memcpy(&id, params->id, sizeof(id));
  } else {
    if (rank == 0) B200_NCCL(nccl, ncclGetUniqueId(&id));
    if (count > 1) {
      B200_RETURN_IF_ERROR(iree_b200_channel_provider_exchange_default_id(
          rank, reinterpret_cast<uint8_t*>(&id)));
    }
  }

  B200_CU(device->cu, cuCtxSetCurrent(device->cu_context));
  ncclConfig_t config = NCCL_CONFIG_INITIALIZER;
  config.blocking = 1;  // nccl_channel.c:113-114
  ncclComm_t comm = nullptr;
  B200_NCCL(nccl, ncclCommInitRankConfig(&comm, count, id, rank, &config));

  auto* channel = new iree_b200_channel_t();
  channel->device = device;
  channel->nccl = nccl;
  channel->comm = comm;
  channel->rank = rank;
  channel->count = count;
  iree_b200_device_retain(device);
  *out_channel = channel;
  return IREE_B200_STATUS_OK;
}

void iree_b200_channel_retain(iree_b200_channel_t* channel) {
  if (channel) channel->ref_count.fetch_add(1);
}

void iree_b200_channel_release(iree_b200_channel_t* channel) {
  if (!channel || channel->ref_count.fetch_sub(1) != 1) return;
  iree_b200_device_t* device = channel->device;
  device->cu->cuCtxSetCurrent(device->cu_context);
  if (channel->comm) channel->nccl->ncclCommDestroy(channel->comm);
  if (channel->symmetric.table) device->cu->cuMemFree(channel->symmetric.table);
  iree_b200_executable_release(channel->symmetric.executable);
  delete channel;
  iree_b200_device_release(device);
}

iree_b200_status_t iree_b200_channel_split(iree_b200_channel_t* base_channel, int32_t color,
                                           int32_t key, uint64_t flags,
                                           iree_b200_channel_t** out_split_channel) {
  (void)flags;
  B200_REQUIRE_ARG(base_channel && out_split_channel);
  *out_split_channel = nullptr;
  const NcclSymbols* nccl = base_channel->nccl;
  iree_b200_device_t* device = base_channel->device;
  B200_CU(device->cu, cuCtxSetCurrent(device->cu_context));
  ncclConfig_t config = NCCL_CONFIG_INITIALIZER;
  config.blocking = 1;
  ncclComm_t split = nullptr;
  B200_NCCL(nccl, ncclCommSplit(base_channel->comm,
                                color == IREE_B200_CHANNEL_NO_COLOR ? NCCL_SPLIT_NOCOLOR : color, key,
                                &split, &config));
  if (color == IREE_B200_CHANNEL_NO_COLOR || !split) return IREE_B200_STATUS_OK;
  int rank = 0, count = 0;
  B200_NCCL(nccl, ncclCommUserRank(split, &rank));
  B200_NCCL(nccl, ncclCommCount(split, &count));
  auto* channel = new iree_b200_channel_t();
  channel->device = device;
  channel->nccl = nccl;
  channel->comm = split;
  channel->rank = rank;
  channel->count = count;
  iree_b200_device_retain(device);
  *out_split_channel = channel;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_channel_attach_symmetric_region(
    iree_b200_channel_t* channel, iree_b200_executable_t* executable,
    const iree_b200_symmetric_region_params_t* params) {
  B200_REQUIRE_ARG(channel && executable && params);
  B200_REQUIRE_ARG(params->peer_bases && params->peer_flag_blocks);
  using Cfg = b200::CollectiveConfig;
  static_assert(Cfg::kMaxWorld == IREE_B200_SYMMETRIC_MAX_RANKS, "header constant out of date");
  static_assert(Cfg::kFlagBytes == IREE_B200_SYMMETRIC_FLAG_BYTES, "header constant out of date");
  if (channel->symmetric.attached) {
    return make_status(IREE_B200_STATUS_ALREADY_EXISTS,
                       "the channel already has a symmetric region attached");
  }
  if (channel->count > Cfg::kMaxWorld) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE,
                       "symmetric regions span at most %d ranks (channel has %d)", Cfg::kMaxWorld,
                       channel->count);
  }
  if (params->max_blocks < 0 || params->max_blocks > Cfg::kMaxBlocks) {
    return make_status(IREE_B200_STATUS_OUT_OF_RANGE, "max_blocks %d out of range [0, %d]",
                       params->max_blocks, Cfg::kMaxBlocks);
  }
  uint64_t table[2 * Cfg::kMaxWorld + 2] = {0};  // + the "range ready" word, zeroed
  for (int r = 0; r < channel->count; ++r) {
    if (!params->peer_bases[r] || !params->peer_flag_blocks[r] || params->peer_bases[r] % 16 ||
        params->peer_flag_blocks[r] % 4) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "symmetric region: mapping of rank %d is null or misaligned", r);
    }
    table[r] = params->peer_bases[r];
    table[Cfg::kMaxWorld + r] = params->peer_flag_blocks[r];
  }
  if (params->multicast_base % 16) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "multicast mapping is misaligned");
  }
  const char* const suffix[3] = {"f16", "bf16", "f32"};
  CUfunction functions[3];
  for (int t = 0; t < 3; ++t) {
    const std::string name = std::string("iree_b200_all_reduce_sum_") + suffix[t] +
                             (params->multicast_base ? "_nvls" : "_p2p");
    int32_t ordinal = -1;
    B200_RETURN_IF_ERROR(iree_b200_executable_lookup_export_by_name(executable, name.c_str(), &ordinal));
    functions[t] = executable->exports[ordinal].function;
  }
  iree_b200_device_t* device = channel->device;
  const CudaSymbols* cu = device->cu;
  B200_CU(cu, cuCtxSetCurrent(device->cu_context));
  CUdeviceptr device_table = 0;
  B200_CU(cu, cuMemAlloc(&device_table, sizeof(table)));
  {
    std::lock_guard<std::mutex> lock(device->mutex);
    CUresult r = cu->cuMemcpyHtoDAsync(device_table, table, sizeof(table), device->dispatch_stream);
    if (r == CUDA_SUCCESS) r = cu->cuStreamSynchronize(device->dispatch_stream);
    if (r != CUDA_SUCCESS) {
      cu->cuMemFree(device_table);
      return cu_status(cu, r, "upload of the symmetric peer table", __FILE__, __LINE__);
    }
  }
  auto& sym = channel->symmetric;
  sym.local_base = (CUdeviceptr)params->peer_bases[channel->rank];
  sym.byte_length = params->byte_length;
  sym.multicast_base = (CUdeviceptr)params->multicast_base;
  sym.table = device_table;
  sym.max_blocks = params->max_blocks ? params->max_blocks : Cfg::kDefaultBlocks;
  sym.executable = executable;
  iree_b200_executable_retain(executable);
  for (int t = 0; t < 3; ++t) sym.all_reduce_sum[t] = functions[t];
  sym.attached = true;
  return IREE_B200_STATUS_OK;
}

void iree_b200_channel_query_collective_counts(const iree_b200_channel_t* channel,
                                               uint64_t* out_symmetric, uint64_t* out_nccl) {
  if (out_symmetric) *out_symmetric = channel ? channel->symmetric_collectives.load() : 0;
  if (out_nccl) *out_nccl = channel ? channel->nccl_collectives.load() : 0;
}

void iree_b200_channel_query_rank_and_count(const iree_b200_channel_t* channel, int32_t* out_rank,
                                            int32_t* out_count) {
  if (out_rank) *out_rank = channel ? channel->rank : 0;
  if (out_count) *out_count = channel ? channel->count : 0;
}

}  // extern "C"
