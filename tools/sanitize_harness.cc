// sanitize_harness.cc — exercises the GPU-less surface of the C-ABI (offline ops object: create /
// select / refuse / release; device creation without a driver; executable sniffing of junk) for
// tools/sanitize_host.sh, which builds the library with ASAN + UBSAN + LSAN. Test infrastructure.
#include <cstdio>
#include <cstring>
#include "iree_b200_hal.h"
#include "iree_b200_ops.h"
int main() {
  for (int i = 0; i < 50; ++i) {
    iree_b200_ops_t* ops = nullptr;
    if (iree_b200_ops_create_offline(nullptr, 0, 148, &ops) != IREE_B200_STATUS_OK) return 1;
    iree_b200_matmul_selection_t sel;
    iree_b200_ops_select_matmul(ops, IREE_B200_ELEMENT_TYPE_BF16, 1, 8192, 8192, 8192, 0, 256, &sel);
    iree_b200_ops_select_matmul(ops, 99, 1, 8, 8, 8, 0, 256, &sel);  // error path
    iree_b200_attention_selection_t asel;
    iree_b200_ops_select_attention(ops, IREE_B200_ELEMENT_TYPE_F16, 32, 4096, 128, 4096, 128, 0.1f, 0, 0, 256, &asel);
    float out[4];
    iree_b200_ops_matmul_host(ops, IREE_B200_ELEMENT_TYPE_BF16, 1, 1, 1, 1, 0, out, out, out);  // refused
    iree_b200_ops_release(ops);
  }
  // No GPU here: device creation must fail cleanly.
  iree_b200_device_params_t p;
  iree_b200_device_params_initialize(&p);
  iree_b200_device_t* dev = nullptr;
  iree_b200_status_t s = iree_b200_device_create(0, &p, &dev);
  printf("device_create -> %s (%s)\n", iree_b200_status_code_string(s), iree_b200_status_message());
  char fmt[16]; size_t sz = 0;
  unsigned char junk[256]; memset(junk, 0x5a, sizeof(junk));
  iree_b200_executable_infer_format(junk, sizeof(junk), sizeof(fmt), fmt, &sz);
  puts("harness done");
  return 0;
}
