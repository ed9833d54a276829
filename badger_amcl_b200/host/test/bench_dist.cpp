// Standalone C++ host of the DISTRIBUTED hot path (no Python, no torch): one process per GPU, each driving the C ABI of
// libbamcl_cuda.so -- bamcl_comm_create_file (NCCL id through a file), bamcl_pf_shard_attach, then per step
// bamcl_pf_shard_update_planar_staged (PlanarScanner::updateSensor over the global set, planar_scanner.cpp:125 ->
// particle_filter.cpp:223) and bamcl_pf_shard_resample (ParticleFilter::updateResample, particle_filter.cpp:423).
// Launch: python -m torch.distributed.run --no-python --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1
//         --master-port P badger_amcl_b200/host/test/bench_dist IN.bin STEPS WARMUP
// (the launcher only provides RANK / LOCAL_RANK / WORLD_SIZE / MASTER_PORT; any other process spawner does).
// <in.bin> is the input format of test_host_api (header, cells, scan, poses of the GLOBAL set). Rank 0 prints one JSON line.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "bamcl.h"

#define CHECK(expr)                                                                          \
  do                                                                                         \
  {                                                                                          \
    if ((expr) != BAMCL_OK)                                                                  \
    {                                                                                        \
      std::fprintf(stderr, "rank %d: %s failed: %s\n", rank, #expr, bamcl_last_error());   \
      return 1;                                                                              \
    }                                                                                        \
  } while (0)

static int env_int(const char* name, int fallback)
{
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : fallback;
}

int main(int argc, char** argv)
{
  const int rank = env_int("RANK", 0), local = env_int("LOCAL_RANK", 0), world = env_int("WORLD_SIZE", 1);
  if (argc != 4)
  {
    std::fprintf(stderr, "usage: bench_dist <in.bin> <steps> <warmup>\n");
    return 2;
  }
  const int steps = std::atoi(argv[2]), warmup = std::atoi(argv[3]);
  std::vector<unsigned char> buf;
  {
    FILE* f = std::fopen(argv[1], "rb");
    if (!f)
      return 2;
    std::fseek(f, 0, SEEK_END);
    buf.resize((size_t)std::ftell(f));
    std::fseek(f, 0, SEEK_SET);
    if (std::fread(buf.data(), 1, buf.size(), f) != buf.size())
      return 2;
    std::fclose(f);
  }
  size_t at = 0;
  auto get = [&](void* dst, size_t bytes) {
    std::memcpy(dst, buf.data() + at, bytes);
    at += bytes;
  };
  int hdr[9];
  double par[3];
  get(hdr, sizeof(hdr));
  get(par, sizeof(par));
  const int size_x = hdr[1], size_y = hdr[2], n = hdr[3], n_beams = hdr[4], min_samples = hdr[5];
  const double resolution = par[0], max_dist = par[1], range_max = par[2];
  std::vector<int8_t> cells((size_t)size_x * size_y);
  get(cells.data(), cells.size());
  std::vector<double> ranges(n_beams), angles(n_beams);
  get(ranges.data(), sizeof(double) * n_beams);
  get(angles.data(), sizeof(double) * n_beams);
  const double* poses = reinterpret_cast<const double*>(buf.data() + at);

  bamcl_ctx* ctx = nullptr;
  CHECK(bamcl_ctx_create(local, &ctx));
  bamcl_comm* comm = nullptr;
  const std::string id_path = std::string("/tmp/bamcl_nccl_id_") + (std::getenv("MASTER_PORT") ? std::getenv("MASTER_PORT") : "0") +
                              "_" + (std::getenv("TORCHELASTIC_RUN_ID") ? std::getenv("TORCHELASTIC_RUN_ID") : "run");
  if (rank == 0)
    std::remove(id_path.c_str());
  CHECK(bamcl_comm_create_file(ctx, id_path.c_str(), rank, world, 120, &comm));
  bamcl_map2d* map = nullptr;
  CHECK(bamcl_map2d_create(ctx, size_x, size_y, resolution, 0.f, 0.f, &map));
  CHECK(bamcl_map2d_set_cells(map, cells.data(), (long)cells.size()));
  CHECK(bamcl_map2d_build_distances(map, max_dist));
  bamcl_pf* pf = nullptr;
  CHECK(bamcl_pf_create(ctx, min_samples, n, &pf));
  CHECK(bamcl_pf_shard_attach(pf, comm));
  bamcl_planar_model model;
  std::memset(&model, 0, sizeof(model));
  model.model_type = BAMCL_PLANAR_LF;
  model.max_beams = n_beams;
  model.z_hit = 0.95;
  model.z_short = 0.1;
  model.z_max = 0.05;
  model.z_rand = 0.05;
  model.sigma_hit = 0.2;
  model.lambda_short = 0.1;
  model.off_map_factor = model.non_free_space_factor = 1.0;
  model.scanner_pose[0] = 0.1;
  CHECK(bamcl_pf_stage_scan(pf, map, &model, ranges.data(), angles.data(), n_beams, range_max));
  long off = 0, cnt = 0;
  CHECK(bamcl_pf_shard_set_count(pf, n, &off, &cnt));
  double* aos = nullptr;
  CHECK(bamcl_host_alloc(sizeof(double) * 4 * (size_t)(cnt > 0 ? cnt : 1), reinterpret_cast<void**>(&aos)));
  for (long i = 0; i < cnt; i++)
  {
    aos[4 * i] = poses[3 * (off + i)];
    aos[4 * i + 1] = poses[3 * (off + i) + 1];
    aos[4 * i + 2] = poses[3 * (off + i) + 2];
    aos[4 * i + 3] = 1.0 / n;
  }
  int valid = 0;
  for (int i = 0; i < n_beams; i++)
    valid += ranges[(size_t)i] < range_max ? 1 : 0;
  double up_ms = 0.0, rs_ms = 0.0;
  bamcl_resample_result res;
  std::memset(&res, 0, sizeof(res));
  for (int it = 0; it < warmup + steps; it++)
  {
    // the same global set every step (untimed): every rank uploads the window it owns, then the set's tree
    CHECK(bamcl_pf_shard_set_count(pf, n, &off, &cnt));
    if (cnt > 0)
      CHECK(bamcl_pf_upload_window_samples(pf, aos, off, cnt));
    bamcl_resample_result init;
    CHECK(bamcl_pf_shard_init_stats(pf, &init));
    CHECK(bamcl_ctx_flush_l2(ctx));
    CHECK(bamcl_ctx_sync(ctx));
    int token = 0;
    CHECK(bamcl_comm_all_reduce_i32(comm, &token, 1, 0));  // barrier
    const auto t0 = std::chrono::steady_clock::now();
    double total = 0.0;
    CHECK(bamcl_pf_shard_update_planar_staged(pf, map, 0, &total));
    const auto t1 = std::chrono::steady_clock::now();
    bamcl_resample_params prm;
    std::memset(&prm, 0, sizeof(prm));
    prm.resample_systematic = 1;
    prm.min_samples = min_samples;
    prm.max_samples = n;
    prm.pop_err = 0.01;
    prm.pop_z = 0.99;
    prm.w_diff = 0.0;
    prm.dist_threshold = 0.5;
    prm.convergence_threshold = 95.0;
    prm.drand48_state = ((uint64_t)12345 << 16) | 0x330E;
    CHECK(bamcl_pf_shard_resample(pf, &prm, &res));
    const auto t2 = std::chrono::steady_clock::now();
    int us[2] = {(int)std::chrono::duration<double, std::micro>(t1 - t0).count(),
                 (int)std::chrono::duration<double, std::micro>(t2 - t1).count()};
    CHECK(bamcl_comm_all_reduce_i32(comm, us, 2, 2));  // max over ranks
    if (it >= warmup)
    {
      up_ms += us[0] * 1e-3;
      rs_ms += us[1] * 1e-3;
    }
  }
  if (rank == 0)
    std::printf("{\"host\": \"C++ (bench_dist.cpp over the C ABI, NCCL id through a file)\", \"n_gpus\": %d, \"global_particles\": %d, "
                "\"beams\": %d, \"valid_beams\": %d, \"steps\": %d, \"update_ms\": %.6f, \"resample_ms\": %.6f, \"ms_per_step\": %.6f, "
                "\"evals_per_s\": %.6e, \"resampled_count\": %d, \"kld_leaf_count\": %d, \"clusters\": %d, "
                "\"timing\": \"host steady_clock around the synchronous calls, max over ranks\"}\n",
                world, n, n_beams, valid, steps, up_ms / steps, rs_ms / steps, (up_ms + rs_ms) / steps,
                (double)n * valid / ((up_ms + rs_ms) / steps * 1e-3), res.sample_count, res.leaf_count, res.cluster_count);
  bamcl_pf_destroy(pf);
  bamcl_map2d_destroy(map);
  bamcl_comm_destroy(comm);
  bamcl_host_free(aos);
  bamcl_ctx_destroy(ctx);
  if (rank == 0)
    std::remove(id_path.c_str());
  return 0;
}
