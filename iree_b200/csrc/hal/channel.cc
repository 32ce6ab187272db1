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

#include <atomic>
#include <chrono>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

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

// The id exchange (stands in for MPI_Bcast, mpi_channel_provider.c:185-190) is a handshake
// through small files next to IREE_B200_CHANNEL_ID_FILE, keyed by a per-process exchange
// sequence number (ranks create their default-id channels in the same order, the assumption
// MPI_Bcast makes too) and by a FRESH random token per rank and exchange:
//   rank r>0  publishes  <path>.<seq>.hello.<r> = token_r            (write + rename)
//   rank 0    waits for every hello, publishes <path>.<seq>.id = {magic, seq, count,
//             token_1..token_{count-1}, id[128]}
//   rank r    accepts the id file only if it carries ITS token, then publishes
//             <path>.<seq>.ack.<r> = token_r
//   rank 0    returns once every ack carries the token it published for that rank — and
//             republishes if a hello changes meanwhile (it had read a stale one) — then
//             removes every file of the exchange.
// A file left behind by an earlier run or an earlier channel can therefore never be taken
// for the current id: it does not carry a token drawn after this exchange started.
namespace {

constexpr uint32_t kIdFileMagic = 0x31444942u;  // "BID1"

bool write_file_atomically(const std::string& path, const void* data, size_t size) {
  const std::string tmp = path + ".tmp." + std::to_string((long)getpid());
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) return false;
  const size_t written = fwrite(data, 1, size, f);
  fclose(f);
  if (written != size || rename(tmp.c_str(), path.c_str()) != 0) {
    unlink(tmp.c_str());
    return false;
  }
  return true;
}

bool read_file_exactly(const std::string& path, void* data, size_t size) {
  struct stat st;
  if (stat(path.c_str(), &st) != 0 || (size_t)st.st_size != size) return false;
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  const size_t got = fread(data, 1, size, f);
  fclose(f);
  return got == size;
}

uint64_t fresh_token() {
  uint64_t token = 0;
  FILE* f = fopen("/dev/urandom", "rb");
  if (f) {
    if (fread(&token, 1, sizeof(token), f) != sizeof(token)) token = 0;
    fclose(f);
  }
  if (!token) {
    token = (uint64_t)std::chrono::steady_clock::now().time_since_epoch().count() ^
            ((uint64_t)getpid() << 32) ^ (uint64_t)(uintptr_t)&token;
  }
  return token | 1;  // never 0
}

std::atomic<uint32_t> g_exchange_sequence{0};

iree_b200_status_t exchange_id(int32_t rank, int32_t count, uint8_t inout_id[IREE_B200_CHANNEL_ID_SIZE]) {
  const char* path = getenv("IREE_B200_CHANNEL_ID_FILE");
  if (!path || !*path) {
    return make_status(IREE_B200_STATUS_FAILED_PRECONDITION,
                       "no channel id given and IREE_B200_CHANNEL_ID_FILE is not set; cannot "
                       "exchange the default id between ranks");
  }
  if (count < 2 || rank < 0 || rank >= count) {
    return make_status(IREE_B200_STATUS_INVALID_ARGUMENT, "id exchange: rank %d of %d", rank, count);
  }
  const uint32_t seq = g_exchange_sequence.fetch_add(1);
  const std::string base = std::string(path) + "." + std::to_string(seq);
  const std::string id_path = base + ".id";
  auto hello_path = [&](int r) { return base + ".hello." + std::to_string(r); };
  auto ack_path = [&](int r) { return base + ".ack." + std::to_string(r); };
  int32_t timeout_s = 120;
  env_int("IREE_B200_CHANNEL_ID_TIMEOUT_S", &timeout_s);
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(timeout_s);
  auto expired = [&] { return std::chrono::steady_clock::now() >= deadline; };
  auto nap = [] { std::this_thread::sleep_for(std::chrono::milliseconds(2)); };
  // {magic, seq, count, reserved} + tokens[count] + id. tokens[0] is rank 0's own nonce for
  // THIS exchange: an acknowledgement must echo it, so that leftover hello + ack files of a
  // dead run (same sequence number, mutually consistent) cannot complete an exchange for ranks
  // that have not arrived yet.
  const size_t payload_size = 16 + 8 * (size_t)count + IREE_B200_CHANNEL_ID_SIZE;
  std::vector<uint8_t> payload(payload_size);

  if (rank != 0) {
    const uint64_t token = fresh_token();
    unlink(ack_path(rank).c_str());
    if (!write_file_atomically(hello_path(rank), &token, sizeof(token))) {
      return make_status(IREE_B200_STATUS_PERMISSION_DENIED, "cannot write %s",
                         hello_path(rank).c_str());
    }
    for (;;) {
      if (read_file_exactly(id_path, payload.data(), payload_size)) {
        uint32_t head[4];
        uint64_t published = 0;
        memcpy(head, payload.data(), sizeof(head));
        memcpy(&published, payload.data() + 16 + 8 * (size_t)rank, sizeof(published));
        if (head[0] == kIdFileMagic && head[1] == seq && head[2] == (uint32_t)count &&
            published == token) {
          memcpy(inout_id, payload.data() + 16 + 8 * (size_t)count, IREE_B200_CHANNEL_ID_SIZE);
          uint64_t ack[2] = {token, 0};
          memcpy(&ack[1], payload.data() + 16, sizeof(uint64_t));  // rank 0's nonce
          if (!write_file_atomically(ack_path(rank), ack, sizeof(ack))) {
            return make_status(IREE_B200_STATUS_PERMISSION_DENIED, "cannot write %s",
                               ack_path(rank).c_str());
          }
          return IREE_B200_STATUS_OK;
        }
      }
      if (expired()) {
        return make_status(IREE_B200_STATUS_DEADLINE_EXCEEDED,
                           "rank %d timed out waiting for channel id #%u at %s", rank, seq,
                           id_path.c_str());
      }
      nap();
    }
  }

  // Rank 0.
  unlink(id_path.c_str());
  std::vector<uint64_t> tokens(count, 0), published(count, 0);
  const uint64_t nonce = fresh_token();
  tokens[0] = nonce;
  for (;;) {
    bool all_hello = true, all_acked = true, changed = false;
    for (int r = 1; r < count; ++r) {
      uint64_t t = 0;
      if (read_file_exactly(hello_path(r), &t, sizeof(t)) && t) {
        if (t != tokens[r]) changed = true;
        tokens[r] = t;
      }
      if (!tokens[r]) all_hello = false;
    }
    if (all_hello && (changed || !published[1])) {
      const uint32_t head[4] = {kIdFileMagic, seq, (uint32_t)count, 0};
      memcpy(payload.data(), head, sizeof(head));
      memcpy(payload.data() + 16, tokens.data(), 8 * (size_t)count);
      memcpy(payload.data() + 16 + 8 * (size_t)count, inout_id, IREE_B200_CHANNEL_ID_SIZE);
      if (!write_file_atomically(id_path, payload.data(), payload_size)) {
        return make_status(IREE_B200_STATUS_DATA_LOSS, "cannot publish channel id at %s",
                           id_path.c_str());
      }
      published = tokens;
    }
    for (int r = 1; r < count; ++r) {
      uint64_t ack[2] = {0, 0};
      if (!published[r] || !read_file_exactly(ack_path(r), ack, sizeof(ack)) ||
          ack[0] != published[r] || ack[1] != nonce) {
        all_acked = false;
      }
    }
    if (all_hello && all_acked) break;
    if (expired()) {
      return make_status(IREE_B200_STATUS_DEADLINE_EXCEEDED,
                         "rank 0 timed out exchanging channel id #%u through %s (%s)", seq,
                         base.c_str(), all_hello ? "acknowledgements missing" : "ranks missing");
    }
    nap();
  }
  unlink(id_path.c_str());
  for (int r = 1; r < count; ++r) {
    unlink(hello_path(r).c_str());
    unlink(ack_path(r).c_str());
  }
  return IREE_B200_STATUS_OK;
}

}  // namespace

iree_b200_status_t iree_b200_channel_provider_exchange_default_id(
    int32_t rank, uint8_t inout_id[IREE_B200_CHANNEL_ID_SIZE]) {
  B200_REQUIRE_ARG(inout_id);
  int32_t default_rank = 0, count = 1;
  B200_RETURN_IF_ERROR(
      iree_b200_channel_provider_query_default_rank_and_count(&default_rank, &count));
  return exchange_id(rank, count, inout_id);
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
      B200_RETURN_IF_ERROR(exchange_id(rank, count, reinterpret_cast<uint8_t*>(&id)));
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
  // bases, flag blocks, then {ready word, status word} (zeroed) and the watchdog limit.
  static_assert(Cfg::kTableBytes % 8 == 0 && Cfg::kWatchdogOffset % 8 == 0, "table layout");
  uint64_t table[Cfg::kTableBytes / 8] = {0};
  int32_t watchdog_ms = 0;
  env_int("IREE_B200_COLLECTIVE_WATCHDOG_MS", &watchdog_ms);
  if (watchdog_ms > 0) {
    int clock_khz = 0;
    const CudaSymbols* cu0 = channel->device->cu;
    cu0->cuDeviceGetAttribute(&clock_khz, CU_DEVICE_ATTRIBUTE_CLOCK_RATE, channel->device->cu_device);
    if (clock_khz <= 0) clock_khz = 2000000;
    table[Cfg::kWatchdogOffset / 8] = (uint64_t)watchdog_ms * (uint64_t)clock_khz;
  }
  int32_t threads = Cfg::kThreads, unroll = Cfg::kUnroll, pipelined = Cfg::kPipelined;
  if (const char* tuning = getenv("IREE_B200_ALL_REDUCE_TUNING")) {
    if (sscanf(tuning, "%d,%d,%d", &threads, &unroll, &pipelined) < 1 || threads < 32 ||
        threads > Cfg::kThreads || threads % 32 ||
        !(unroll == 1 || unroll == 2 || unroll == 4 || unroll == 8) || (pipelined && unroll == 8)) {
      return make_status(IREE_B200_STATUS_INVALID_ARGUMENT,
                         "IREE_B200_ALL_REDUCE_TUNING=\"%s\": want threads,unroll,pipelined", tuning);
    }
  }
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
  sym.threads = threads;
  sym.tuning = (uint32_t)unroll | (pipelined ? 0x100u : 0u);
  sym.executable = executable;
  iree_b200_executable_retain(executable);
  for (int t = 0; t < 3; ++t) sym.all_reduce_sum[t] = functions[t];
  sym.attached = true;
  return IREE_B200_STATUS_OK;
}

iree_b200_status_t iree_b200_channel_query_collective_status(iree_b200_channel_t* channel,
                                                             uint32_t* out_status) {
  B200_REQUIRE_ARG(channel && out_status);
  *out_status = 0;
  if (!channel->symmetric.attached) return IREE_B200_STATUS_OK;
  const CudaSymbols* cu = channel->device->cu;
  B200_CU(cu, cuCtxSetCurrent(channel->device->cu_context));
  // On a stream of its own: the dispatch stream may be busy behind the very kernel asked about.
  CUstream stream = nullptr;
  B200_CU(cu, cuStreamCreate(&stream, CU_STREAM_NON_BLOCKING));
  CUresult r = cu->cuMemcpyDtoHAsync(
      out_status, channel->symmetric.table + b200::CollectiveConfig::kStatusOffset, sizeof(uint32_t),
      stream);
  if (r == CUDA_SUCCESS) r = cu->cuStreamSynchronize(stream);
  cu->cuStreamDestroy(stream);
  if (r != CUDA_SUCCESS) return cu_status(cu, r, "read of the collective status word", __FILE__, __LINE__);
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
