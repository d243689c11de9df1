// Link check: the host shim with the real ABI table. Running it needs a GPU (it creates a handle); building it does not.
#include "lidar_gpu.hpp"
#include <cstdio>
int main(int argc, char **) {
  mins_gpu::ml_api api = mins_gpu::ml_api::real();
  std::printf("ml_api bound: %d entries\n", (int)(sizeof(api) / sizeof(void *)));
  if (argc > 1) { // any argument: also create and destroy a handle (GPU box only)
    ml_config cfg{0, 1024, 4096, 0, 0.f};
    ml_handle *h = nullptr;
    int rc = api.create(&cfg, &h);
    std::printf("ml_create -> %d\n", rc);
    if (rc == ML_OK)
      api.destroy(h);
    return rc == ML_OK ? 0 : 1;
  }
  return 0;
}
