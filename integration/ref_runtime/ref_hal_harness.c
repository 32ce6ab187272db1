// ref_hal_harness — runs ONE dispatch of a CDA1 executable through the REFERENCE runtime
// (iree_hal_* API, the reference's own CUDA HAL driver), reads the result bindings back and
// times a batch. Test infrastructure: the drop-in check of SURVEY.md §8(b).
//
// It does what the reference's tools/iree-benchmark-executable-main.c:191-300 does —
// executable cache -> command buffer dispatch -> queue_execute -> semaphore wait — plus the
// two things that tool cannot do: load raw operand bytes from files and write the output
// bindings back to files, so a test can compare them with the oracle.
//
//   ref_hal_harness --device=cuda --executable_file=k.cda1 --entry_point=N
//       --workgroup_count=x,y,z --constant=u32 ... --binding=<bytes>[=@file] ...
//       --output=<binding index>=<file> ... [--batch_size=B --iterations=I]
//
// Built against /root/reference by tools/build_ref_runtime.sh (never part of the product).

#include <inttypes.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "iree/async/frontier_tracker.h"
#include "iree/async/util/proactor_pool.h"
#include "iree/base/api.h"
#include "iree/base/threading/numa.h"
#include "iree/base/tooling/flags.h"
#include "iree/hal/api.h"
#include "iree/io/file_contents.h"
#include "iree/tooling/device_util.h"

IREE_FLAG(string, executable_format, "PTXE", "Executable format (CDA1 containers are PTXE).");
IREE_FLAG(string, executable_file, "", "Path of the CDA1 executable.");
IREE_FLAG(int32_t, entry_point, 0, "Export ordinal.");
IREE_FLAG(string, workgroup_count, "1,1,1", "x,y,z");
IREE_FLAG(int32_t, batch_size, 1, "Dispatches (each followed by a barrier) per submission.");
IREE_FLAG(int32_t, iterations, 0, "Timed submissions after the checked one (0 = none).");
IREE_FLAG_LIST(string, constant, "uint32 push constant (repeatable, in order).");
IREE_FLAG_LIST(string, binding, "<byte length>[=@file] (repeatable, in order).");
IREE_FLAG_LIST(string, output, "<binding index>=<file>: written after the checked run.");

#define HARNESS_MAX 64

static iree_status_t parse_u64(iree_string_view_t s, uint64_t* out) {
  if (!iree_string_view_atoi_uint64(s, out)) {
    return iree_make_status(IREE_STATUS_INVALID_ARGUMENT, "not an integer: `%.*s`", (int)s.size,
                            s.data);
  }
  return iree_ok_status();
}

static iree_status_t submit_and_wait(iree_hal_device_t* device, iree_hal_command_buffer_t* cb,
                                     iree_hal_semaphore_t* fence, uint64_t* fence_value) {
  ++*fence_value;
  iree_hal_semaphore_list_t signals = {
      .count = 1, .semaphores = &fence, .payload_values = fence_value};
  IREE_RETURN_IF_ERROR(iree_hal_device_queue_execute(
      device, IREE_HAL_QUEUE_AFFINITY_ANY, iree_hal_semaphore_list_empty(), signals, cb,
      iree_hal_buffer_binding_table_empty(), IREE_HAL_EXECUTE_FLAG_NONE));
  return iree_hal_semaphore_wait(fence, *fence_value, iree_infinite_timeout(),
                                 IREE_ASYNC_WAIT_FLAG_NONE);
}

static iree_status_t run(iree_allocator_t host_allocator) {
  // Device through the reference's driver registry (--device=cuda).
  iree_async_proactor_pool_t* proactor_pool = NULL;
  IREE_RETURN_IF_ERROR(iree_async_proactor_pool_create(
      iree_numa_node_count(), NULL, iree_async_proactor_pool_options_default(), host_allocator,
      &proactor_pool));
  iree_hal_device_create_params_t create_params = iree_hal_device_create_params_default();
  create_params.proactor_pool = proactor_pool;
  iree_hal_device_t* device = NULL;
  iree_status_t status = iree_hal_create_device_from_flags(
      iree_hal_available_driver_registry(), iree_hal_default_device_uri(), &create_params,
      host_allocator, &device);
  iree_async_proactor_pool_release(proactor_pool);
  IREE_RETURN_IF_ERROR(status);
  iree_async_frontier_tracker_t* tracker = NULL;
  IREE_RETURN_IF_ERROR(iree_async_frontier_tracker_create(
      iree_async_frontier_tracker_options_default(), host_allocator, &tracker));
  iree_hal_device_group_t* group = NULL;
  status = iree_hal_device_group_create_from_device(device, tracker, host_allocator, &group);
  iree_async_frontier_tracker_release(tracker);
  IREE_RETURN_IF_ERROR(status);

  // Executable: the reference loader verifies and loads the CDA1 container.
  iree_hal_executable_cache_t* cache = NULL;
  IREE_RETURN_IF_ERROR(iree_hal_executable_cache_create(device, IREE_SV("harness"), &cache));
  iree_io_file_contents_t* image = NULL;
  IREE_RETURN_IF_ERROR(iree_io_file_contents_read(iree_make_cstring_view(FLAG_executable_file),
                                                  host_allocator, &image));
  iree_hal_executable_params_t exe_params;
  iree_hal_executable_params_initialize(&exe_params);
  exe_params.caching_mode = IREE_HAL_EXECUTABLE_CACHING_MODE_ALLOW_OPTIMIZATION |
                            IREE_HAL_EXECUTABLE_CACHING_MODE_ALIAS_PROVIDED_DATA;
  exe_params.executable_format = iree_make_cstring_view(FLAG_executable_format);
  exe_params.executable_data = image->const_buffer;
  iree_hal_executable_t* executable = NULL;
  IREE_RETURN_IF_ERROR(
      iree_hal_executable_cache_prepare_executable(cache, &exe_params, &executable));
  fprintf(stdout, "{\"loaded\": true, \"function_count\": %" PRIhsz "}\n",
          iree_hal_executable_function_count(executable));

  // Bindings: device-local buffers filled from files (or zero).
  const iree_flag_string_list_t binding_flags = FLAG_binding_list();
  if (binding_flags.count > HARNESS_MAX) {
    return iree_make_status(IREE_STATUS_OUT_OF_RANGE, "too many bindings");
  }
  iree_hal_buffer_t* buffers[HARNESS_MAX] = {0};
  uint64_t lengths[HARNESS_MAX] = {0};
  iree_hal_buffer_ref_t refs[HARNESS_MAX];
  for (iree_host_size_t i = 0; i < binding_flags.count; ++i) {
    iree_string_view_t size_str, file_str;
    iree_string_view_split(binding_flags.values[i], '=', &size_str, &file_str);
    IREE_RETURN_IF_ERROR(parse_u64(size_str, &lengths[i]));
    iree_hal_buffer_params_t params = {0};
    params.type = IREE_HAL_MEMORY_TYPE_DEVICE_LOCAL;
    params.usage = IREE_HAL_BUFFER_USAGE_DISPATCH_STORAGE | IREE_HAL_BUFFER_USAGE_TRANSFER;
    IREE_RETURN_IF_ERROR(iree_hal_allocator_allocate_buffer(
        iree_hal_device_allocator(device), params, lengths[i] ? lengths[i] : 16, &buffers[i]));
    if (lengths[i]) {
      void* host = NULL;
      IREE_RETURN_IF_ERROR(iree_allocator_malloc(host_allocator, lengths[i], &host));
      memset(host, 0, lengths[i]);
      if (iree_string_view_starts_with(file_str, IREE_SV("@"))) {
        char path[1024];
        snprintf(path, sizeof(path), "%.*s", (int)file_str.size - 1, file_str.data + 1);
        FILE* f = fopen(path, "rb");
        if (!f || fread(host, 1, lengths[i], f) != lengths[i]) {
          return iree_make_status(IREE_STATUS_NOT_FOUND, "cannot read %" PRIu64 " bytes of %s",
                                  lengths[i], path);
        }
        fclose(f);
      }
      IREE_RETURN_IF_ERROR(iree_hal_device_transfer_h2d(
          device, host, buffers[i], 0, lengths[i], IREE_HAL_TRANSFER_BUFFER_FLAG_DEFAULT,
          iree_infinite_timeout()));
      iree_allocator_free(host_allocator, host);
    }
    refs[i] = iree_hal_make_buffer_ref(buffers[i], 0, IREE_HAL_WHOLE_BUFFER);
  }

  const iree_flag_string_list_t constant_flags = FLAG_constant_list();
  uint32_t constants[HARNESS_MAX] = {0};
  if (constant_flags.count > HARNESS_MAX) {
    return iree_make_status(IREE_STATUS_OUT_OF_RANGE, "too many constants");
  }
  for (iree_host_size_t i = 0; i < constant_flags.count; ++i) {
    uint64_t v = 0;
    IREE_RETURN_IF_ERROR(parse_u64(constant_flags.values[i], &v));
    constants[i] = (uint32_t)v;
  }
  uint32_t count[3] = {1, 1, 1};
  if (sscanf(FLAG_workgroup_count, "%u,%u,%u", &count[0], &count[1], &count[2]) != 3) {
    return iree_make_status(IREE_STATUS_INVALID_ARGUMENT, "--workgroup_count=x,y,z");
  }

  // One command buffer with |batch_size| x (dispatch + barrier), as the reference tool records.
  iree_hal_command_buffer_t* cb = NULL;
  IREE_RETURN_IF_ERROR(iree_hal_command_buffer_create(
      device, IREE_HAL_COMMAND_BUFFER_MODE_DEFAULT, IREE_HAL_COMMAND_CATEGORY_DISPATCH,
      IREE_HAL_QUEUE_AFFINITY_ANY, 0, &cb));
  IREE_RETURN_IF_ERROR(iree_hal_command_buffer_begin(cb));
  iree_hal_buffer_ref_list_t binding_list = {.count = binding_flags.count, .values = refs};
  iree_hal_dispatch_config_t config =
      iree_hal_make_static_dispatch_config(count[0], count[1], count[2]);
  for (int32_t i = 0; i < FLAG_batch_size; ++i) {
    IREE_RETURN_IF_ERROR(iree_hal_command_buffer_dispatch(
        cb, executable, iree_hal_executable_function_from_index((uint32_t)FLAG_entry_point),
        config, iree_make_const_byte_span(constants, constant_flags.count * sizeof(uint32_t)),
        binding_list, IREE_HAL_DISPATCH_FLAG_NONE));
    IREE_RETURN_IF_ERROR(iree_hal_command_buffer_execution_barrier(
        cb, IREE_HAL_EXECUTION_STAGE_COMMAND_RETIRE, IREE_HAL_EXECUTION_STAGE_COMMAND_ISSUE,
        IREE_HAL_EXECUTION_BARRIER_FLAG_NONE, 0, NULL, 0, NULL));
  }
  IREE_RETURN_IF_ERROR(iree_hal_command_buffer_end(cb));

  iree_hal_semaphore_t* fence = NULL;
  uint64_t fence_value = 0;
  IREE_RETURN_IF_ERROR(iree_hal_semaphore_create(device, IREE_HAL_QUEUE_AFFINITY_ANY, 0,
                                                 IREE_HAL_SEMAPHORE_FLAG_DEFAULT, &fence));
  IREE_RETURN_IF_ERROR(submit_and_wait(device, cb, fence, &fence_value));

  // Read the requested bindings back.
  const iree_flag_string_list_t output_flags = FLAG_output_list();
  for (iree_host_size_t i = 0; i < output_flags.count; ++i) {
    iree_string_view_t index_str, file_str;
    iree_string_view_split(output_flags.values[i], '=', &index_str, &file_str);
    uint64_t index = 0;
    IREE_RETURN_IF_ERROR(parse_u64(index_str, &index));
    if (index >= binding_flags.count) {
      return iree_make_status(IREE_STATUS_OUT_OF_RANGE, "--output binding %" PRIu64, index);
    }
    void* host = NULL;
    IREE_RETURN_IF_ERROR(iree_allocator_malloc(host_allocator, lengths[index], &host));
    IREE_RETURN_IF_ERROR(iree_hal_device_transfer_d2h(
        device, buffers[index], 0, host, lengths[index], IREE_HAL_TRANSFER_BUFFER_FLAG_DEFAULT,
        iree_infinite_timeout()));
    char path[1024];
    snprintf(path, sizeof(path), "%.*s", (int)file_str.size, file_str.data);
    FILE* f = fopen(path, "wb");
    if (!f || fwrite(host, 1, lengths[index], f) != lengths[index]) {
      return iree_make_status(IREE_STATUS_PERMISSION_DENIED, "cannot write %s", path);
    }
    fclose(f);
    iree_allocator_free(host_allocator, host);
  }

  // Timing: wall clock around submissions of the whole batch (what iree-benchmark-executable
  // reports per dispatch, submission overhead included).
  if (FLAG_iterations > 0) {
    for (int i = 0; i < 3; ++i) {
      IREE_RETURN_IF_ERROR(submit_and_wait(device, cb, fence, &fence_value));
    }
    iree_time_t t0 = iree_time_now();
    for (int32_t i = 0; i < FLAG_iterations; ++i) {
      IREE_RETURN_IF_ERROR(submit_and_wait(device, cb, fence, &fence_value));
    }
    double ns = (double)(iree_time_now() - t0);
    fprintf(stdout,
            "{\"ms_per_dispatch\": %.6f, \"batch_size\": %d, \"iterations\": %d, "
            "\"timing\": \"host wall clock around queue_execute + semaphore wait\"}\n",
            ns / 1e6 / ((double)FLAG_iterations * FLAG_batch_size), FLAG_batch_size,
            FLAG_iterations);
  }

  iree_hal_semaphore_release(fence);
  iree_hal_command_buffer_release(cb);
  for (iree_host_size_t i = 0; i < binding_flags.count; ++i) iree_hal_buffer_release(buffers[i]);
  iree_hal_executable_release(executable);
  iree_io_file_contents_free(image);
  iree_hal_executable_cache_release(cache);
  iree_hal_device_group_release(group);
  iree_hal_device_release(device);
  return iree_ok_status();
}

int main(int argc, char** argv) {
  iree_flags_set_usage("ref_hal_harness",
                       "One dispatch of a CDA1 executable through the reference runtime.\n");
  iree_flags_parse_checked(IREE_FLAGS_PARSE_MODE_DEFAULT, &argc, &argv);
  iree_status_t status = run(iree_allocator_system());
  if (!iree_status_is_ok(status)) {
    iree_status_fprint(stderr, status);
    iree_status_free(status);
    return EXIT_FAILURE;
  }
  return EXIT_SUCCESS;
}
