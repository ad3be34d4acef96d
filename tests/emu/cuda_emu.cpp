// TEST INFRASTRUCTURE — see cuda_emu.h.
#include "cuda_emu.h"

thread_local uint3_emu threadIdx, blockIdx;
dim3 blockDim, gridDim;

namespace emu {
Barrier g_block_bar;
std::vector<WarpBox>* g_warps = nullptr;
thread_local int t_linear = 0;
unsigned char* g_dyn_smem = nullptr;

namespace {
struct Pool {
  std::vector<std::thread> th;
  std::mutex m;
  std::condition_variable cv_go, cv_done;
  unsigned gen = 0;
  int n_active = 0, n_done = 0;
  bool quit = false;
  std::function<void(int)> job;
  void ensure(int n) {
    while ((int)th.size() < n) {
      int id = (int)th.size();
      th.emplace_back([this, id] { loop(id); });
    }
  }
  void loop(int id) {
    unsigned seen = 0;
    for (;;) {
      std::function<void(int)> j;
      {
        std::unique_lock<std::mutex> lk(m);
        cv_go.wait(lk, [&] { return quit || (gen != seen); });
        if (quit) return;
        seen = gen;
        if (id >= n_active) continue;
        j = job;
      }
      j(id);
      {
        std::unique_lock<std::mutex> lk(m);
        if (++n_done == n_active) cv_done.notify_all();
      }
    }
  }
  void run(int n, std::function<void(int)> f) {
    ensure(n);
    std::unique_lock<std::mutex> lk(m);
    job = std::move(f);
    n_active = n;
    n_done = 0;
    ++gen;
    cv_go.notify_all();
    cv_done.wait(lk, [&] { return n_done == n_active; });
  }
  ~Pool() {
    {
      std::unique_lock<std::mutex> lk(m);
      quit = true;
      cv_go.notify_all();
    }
    for (auto& t : th) t.join();
  }
};
Pool g_pool;
}  // namespace

void launch_impl(dim3 grid, dim3 block, size_t dyn_smem, const std::function<void()>& body) {
  const int nthreads = (int)(block.x * block.y * block.z);
  const int nwarps = (nthreads + 31) / 32;
  blockDim = block;
  gridDim = grid;
  std::vector<unsigned char> dyn(dyn_smem + 1024);
  // 1024-byte aligned like the real dynamic shared window
  g_dyn_smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(dyn.data()) + 1023) & ~uintptr_t(1023));
  std::vector<WarpBox> warps(nwarps);
  g_warps = &warps;
  for (unsigned bz = 0; bz < grid.z; ++bz)
    for (unsigned by = 0; by < grid.y; ++by)
      for (unsigned bx = 0; bx < grid.x; ++bx) {
        g_block_bar.reset(nthreads);
        for (int w = 0; w < nwarps; ++w) warps[w].bar.reset(std::min(32, nthreads - 32 * w));
        g_pool.run(nthreads, [&](int tid) {
          t_linear = tid;
          threadIdx.x = tid % block.x;
          threadIdx.y = (tid / block.x) % block.y;
          threadIdx.z = tid / (block.x * block.y);
          blockIdx.x = bx;
          blockIdx.y = by;
          blockIdx.z = bz;
          body();
        });
      }
  g_warps = nullptr;
  g_dyn_smem = nullptr;
}
}  // namespace emu
