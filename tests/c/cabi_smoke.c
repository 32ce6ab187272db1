/* The drop-in boundary from plain C11 (the reference's host language): both headers compile
 * as C with -Wall -Wextra -Werror -pedantic, the library links, and — on a machine without a
 * GPU — every compute entry point fails with a status instead of falling back to the CPU.
 * Built and run by tests/test_cabi.py; prints one line per check and exits non-zero on the
 * first surprise. */
#include <stdio.h>
#include <string.h>

#include "iree_b200_hal.h"
#include "iree_b200_ops.h"
#include "iree_b200_tma_recipe.h"

#define CHECK(cond)                                         \
  do {                                                      \
    if (!(cond)) {                                          \
      printf("FAILED line %d: %s\n", __LINE__, #cond);      \
      return 1;                                             \
    }                                                       \
  } while (0)

int main(int argc, char** argv) {
  const int expect_gpu = argc > 1 && strcmp(argv[1], "gpu") == 0;

  /* Status codes are iree_status_code_t numbers. */
  CHECK(strcmp(iree_b200_status_code_string(IREE_B200_STATUS_OK), "OK") == 0);
  CHECK(strcmp(iree_b200_status_code_string(IREE_B200_STATUS_RESOURCE_EXHAUSTED),
               "RESOURCE_EXHAUSTED") == 0);

  /* The builtin executable is a CDA1 container. */
  size_t size = 0;
  const unsigned char* blob = (const unsigned char*)iree_b200_ops_builtin_executable_data(&size);
  CHECK(blob && size > 72 && memcmp(blob, "CDA1", 4) == 0 && memcmp(blob + 68, "CDA1", 4) == 0);

  /* Selection without a device: the arithmetic of the transform spec's count regions. */
  iree_b200_ops_t* offline = NULL;
  CHECK(iree_b200_ops_create_offline(NULL, 0, /*sm_count=*/148, &offline) == IREE_B200_STATUS_OK);
  iree_b200_matmul_selection_t sel;
  CHECK(iree_b200_ops_select_matmul(offline, IREE_B200_ELEMENT_TYPE_BF16, 1, 8192, 8192, 8192,
                                    IREE_B200_MATMUL_FLAG_DEVICE_MAPS, 256, &sel) ==
        IREE_B200_STATUS_OK);
  CHECK(sel.workgroup_count[0] == 148 && sel.constants[0] == 8192);
  /* ... and it never computes. */
  CHECK(iree_b200_ops_matmul_host(offline, IREE_B200_ELEMENT_TYPE_BF16, 1, 8, 8, 8, 0, blob, blob,
                                  (void*)blob) == IREE_B200_STATUS_FAILED_PRECONDITION);
  CHECK(strlen(iree_b200_status_message()) > 0);
  iree_b200_ops_release(offline);

  /* The no-allocation entry point validates before it touches a device. */
  CHECK(iree_b200_attention_decode_no_alloc(0, IREE_B200_ELEMENT_TYPE_F16, 9000, 1, 1024, 0.1f, blob,
                                            blob, blob, (void*)blob, NULL) ==
        IREE_B200_STATUS_RESOURCE_EXHAUSTED);
  CHECK(strstr(iree_b200_status_message(), "IREE_B200_NO_ALLOC_SCRATCH_BYTES") != NULL);
  CHECK(iree_b200_attention_decode_no_alloc(0, IREE_B200_ELEMENT_TYPE_F32, 4, 1, 1024, 0.1f, blob,
                                            blob, blob, (void*)blob, NULL) ==
        IREE_B200_STATUS_INVALID_ARGUMENT);

  /* A device: present on the GPU box, UNAVAILABLE (loudly) elsewhere. */
  iree_b200_device_params_t params;
  iree_b200_device_params_initialize(&params);
  iree_b200_device_t* device = NULL;
  const iree_b200_status_t status = iree_b200_device_create(0, &params, &device);
  if (expect_gpu) {
    CHECK(status == IREE_B200_STATUS_OK && device);
    iree_b200_device_release(device);
  } else {
    CHECK(status == IREE_B200_STATUS_UNAVAILABLE && !device);
  }
  CHECK(sizeof(iree_b200_tma_recipe_t) == 8 + IREE_B200_TMA_MAX_MAPS * sizeof(iree_b200_tma_map_t));
  printf("cabi_smoke ok (%s)\n", expect_gpu ? "gpu" : "no gpu");
  return 0;
}
