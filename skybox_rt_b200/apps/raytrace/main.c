/*
 * raytrace -- the RT regression app, host side, plain C over the vx_* API (include/vortex.h).
 *
 * Written in the idiom of the reference's graphics apps, step for step
 * (tests/regression/draw3d/main.cpp:392-517): parse_args -> vx_dev_open -> caps check -> host
 * pre-pass (draw3d: graphics::Binning, main.cpp:190; here: soup -> BVH2) -> vx_mem_alloc +
 * vx_copy_to_dev per buffer (:198-211) -> kernel_arg with device addresses -> vx_upload_bytes
 * (:346) -> vx_start (:353) -> vx_ready_wait (:357) -> vx_copy_from_dev of the colour buffer and
 * bottom-up save (:383-386) -> compare against a reference image, exit code = #errors (:505-514).
 * Every call is wrapped in RT_CHECK (main.cpp:23-31).  The reference has no RT app (SURVEY.md
 * section 0); what is reused is this host idiom and the framebuffer conventions.
 *
 * Multi-GPU (-g N): one process per GPU, the scene replicated, image rows dealt to the ranks in blocks of 8
 * (kernel_arg_t.rank_id / num_ranks / rows_per_block), and one vx_cuda_frame_gather (NCCL over NVLink, enqueued
 * behind the frame on the launch stream) brings the strips to rank 0, which saves and compares.  Started without
 * RANK / WORLD_SIZE in the environment the program forks its N ranks itself; under a launcher that sets them
 * (torchrun, mpirun -x) every process is one rank.  The NCCL id travels through a file (SRT_COMM_ID_FILE).
 */
#include <getopt.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include "png_io.h"
#include "skybox_rt.h"
#include "vortex.h"
#include "vortex_cuda.h"

#define RT_CHECK(_expr)                                         \
  do {                                                          \
    int _ret = _expr;                                           \
    if (0 == _ret) break;                                       \
    printf("Error: '%s' returned %d!\n", #_expr, (int)_ret);    \
    cleanup();                                                  \
    exit(-1);                                                   \
  } while (0)

static const char* output_file = "output.png";
static const char* reference_file = NULL;
static const char* scene_name = "cornell";
static uint32_t dst_width = 64, dst_height = 64, spp = 1, mode = SRT_MODE_BOUNCE, seed = 1234;
static uint32_t quads_x = 250, quads_y = 200;
static uint32_t num_gpus = 1, rank_id = 0;
#define ROWS_PER_BLOCK 8

static vx_device_h device = NULL;
static vx_buffer_h krnl_buffer = NULL, args_buffer = NULL, node_buffer = NULL, tri_buffer = NULL, color_buffer = NULL;
static vx_buffer_h frame_buffer = NULL; /* rank 0 of a multi-GPU run: the assembled frame */
static vx_comm_h comm = NULL;
static srt_triangle_t* soup = NULL;
static srt_bvh2_t bvh;
static uint32_t* host_fb = NULL;
static kernel_arg_t kernel_arg;

static void show_usage(void) {
  printf("Raytrace Test.\n");
  printf("Usage: [-s scene: cornell|terrain] [-n quads_x] [-q quads_y] [-w width] [-h height] [-m mode 0..2]\n"
         "       [-p spp] [-e seed] [-o output.png] [-r reference.png] [-g gpus]\n");
}

static void parse_args(int argc, char** argv) {
  int c;
  while ((c = getopt(argc, argv, "s:n:q:w:h:m:p:e:o:r:g:?")) != -1) {
    switch (c) {
      case 's': scene_name = optarg; break;
      case 'n': quads_x = (uint32_t)atoi(optarg); break;
      case 'q': quads_y = (uint32_t)atoi(optarg); break;
      case 'w': dst_width = (uint32_t)atoi(optarg); break;
      case 'h': dst_height = (uint32_t)atoi(optarg); break;
      case 'm': mode = (uint32_t)atoi(optarg); break;
      case 'p': spp = (uint32_t)atoi(optarg); break;
      case 'e': seed = (uint32_t)atoi(optarg); break;
      case 'o': output_file = optarg; break;
      case 'r': reference_file = optarg; break;
      case 'g': num_gpus = (uint32_t)atoi(optarg); if (num_gpus < 1 || num_gpus > 64) { show_usage(); exit(-1); } break;
      default: show_usage(); exit(c == '?' ? 0 : -1);
    }
  }
}

static void cleanup(void) {
  if (device) {
    vx_cuda_comm_destroy(comm); comm = NULL;
    vx_mem_free(frame_buffer);
    vx_mem_free(krnl_buffer);
    vx_mem_free(args_buffer);
    vx_mem_free(node_buffer);
    vx_mem_free(tri_buffer);
    vx_mem_free(color_buffer);
    vx_dev_close(device);
    device = NULL;
  }
  free(soup); soup = NULL;
  srt_bvh2_free(&bvh);
  free(host_fb); host_fb = NULL;
}

static double now_ms(void) {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return (double)t.tv_sec * 1e3 + (double)t.tv_nsec * 1e-6;
}

/* rank 0 makes the NCCL id and publishes it through a file (written under a temporary name, then renamed);
 * the other ranks wait for the file. */
static int exchange_comm_id(const char* path, void* id) {
  char tmp[600];
  FILE* f;
  int tries;
  if (rank_id == 0) {
    if (vx_cuda_comm_unique_id(id) != 0) return -1;
    snprintf(tmp, sizeof(tmp), "%s.tmp", path);
    f = fopen(tmp, "wb");
    if (!f) return -1;
    if (fwrite(id, 1, VX_CUDA_COMM_ID_BYTES, f) != VX_CUDA_COMM_ID_BYTES) { fclose(f); return -1; }
    fclose(f);
    return rename(tmp, path);
  }
  for (tries = 0; tries < 6000; ++tries) { /* up to 60 s */
    f = fopen(path, "rb");
    if (f) {
      size_t n = fread(id, 1, VX_CUDA_COMM_ID_BYTES, f);
      fclose(f);
      if (n == VX_CUDA_COMM_ID_BYTES) return 0;
    }
    usleep(10000);
  }
  return -1;
}

/* -g N without a launcher: fork the N ranks (before anything touches CUDA) and wait for them */
static int fork_ranks(int argc, char** argv) {
  char id_file[256], buf[32];
  pid_t pids[64];
  uint32_t r;
  int status, rc = 0;
  (void)argc;
  snprintf(id_file, sizeof(id_file), "/tmp/srt_comm_id.%ld.%ld", (long)getpid(), (long)time(NULL));
  setenv("SRT_COMM_ID_FILE", id_file, 1);
  snprintf(buf, sizeof(buf), "%u", num_gpus);
  setenv("WORLD_SIZE", buf, 1);
  for (r = 0; r < num_gpus; ++r) {
    pid_t pid = fork();
    if (pid < 0) return -1;
    if (pid == 0) {
      snprintf(buf, sizeof(buf), "%u", r);
      setenv("RANK", buf, 1);
      setenv("LOCAL_RANK", buf, 1);
      execv("/proc/self/exe", argv);
      _exit(127);
    }
    pids[r] = pid;
  }
  for (r = 0; r < num_gpus; ++r) {
    pid_t done = wait(&status);
    uint32_t k, who = num_gpus;
    if (done < 0) { rc = -1; continue; }
    for (k = 0; k < num_gpus; ++k)
      if (pids[k] == done) { pids[k] = 0; who = k; }
    if (WIFEXITED(status) && WEXITSTATUS(status) == 0) continue;
    /* ranks other than 0 are silent (their stdout is /dev/null), so the parent says which one failed */
    if (!rc) {
      rc = WIFEXITED(status) ? WEXITSTATUS(status) : -1;
      printf("Error: rank %u of %u failed (%s %d)\n", who, num_gpus, WIFEXITED(status) ? "exit code" : "signal",
             WIFEXITED(status) ? WEXITSTATUS(status) : WTERMSIG(status));
    }
    /* a rank that died leaves the others waiting in a collective: end exactly the processes started here */
    for (k = 0; k < num_gpus; ++k)
      if (pids[k] > 0) kill(pids[k], SIGTERM);
  }
  unlink(id_file);
  return rc;
}

int main(int argc, char* argv[]) {
  uint64_t isa_flags, num_cores, num_warps, num_threads, addr, v;
  uint32_t num_tris = 0, cbuf_size, strip_rows, frame_size;
  srt_kernel_image_t image;
  double t0, t1;

  parse_args(argc, argv);
  if (num_gpus > 1) {
    const char* env_rank = getenv("RANK");
    const char* env_world = getenv("WORLD_SIZE");
    const char* env_local = getenv("LOCAL_RANK");
    if (!env_rank || !env_world) return fork_ranks(argc, argv);
    if ((uint32_t)atoi(env_world) != num_gpus) {
      printf("Error: -g %u but WORLD_SIZE=%s\n", num_gpus, env_world);
      return -1;
    }
    rank_id = (uint32_t)atoi(env_rank);
    setenv("VX_CUDA_DEVICE", env_local ? env_local : env_rank, 1); /* which GPU vx_dev_open takes */
    if (rank_id != 0) { /* one voice on the terminal */
      if (!freopen("/dev/null", "w", stdout)) return -1;
    }
  }
  memset(&bvh, 0, sizeof(bvh));
  memset(&kernel_arg, 0, sizeof(kernel_arg));

  printf("open device connection\n");
  RT_CHECK(vx_dev_open(&device));
  RT_CHECK(vx_dev_caps(device, VX_CAPS_ISA_FLAGS, &isa_flags));
  if (0 == (isa_flags & VX_ISA_EXT_RT)) {
    printf("RT ISA extension is needed!\n");
    cleanup();
    return -1;
  }
  RT_CHECK(vx_dev_caps(device, VX_CAPS_NUM_CORES, &num_cores));
  RT_CHECK(vx_dev_caps(device, VX_CAPS_NUM_WARPS, &num_warps));
  RT_CHECK(vx_dev_caps(device, VX_CAPS_NUM_THREADS, &num_threads));
  printf("number of tasks: %llu\n", (unsigned long long)(num_cores * num_warps * num_threads));

  /* host pre-pass: scene -> BVH2 + triangle records (the role of graphics::Binning) */
  printf("build scene '%s'\n", scene_name);
  if (!strcmp(scene_name, "terrain")) {
    RT_CHECK(srt_scene_terrain(quads_x, quads_y, 0x5EED5EEDu, &soup, &num_tris));
    srt_camera_terrain(&kernel_arg);
  } else {
    RT_CHECK(srt_scene_cornell(&soup, &num_tris));
    srt_camera_cornell(&kernel_arg);
  }
  t0 = now_ms();
  RT_CHECK(srt_bvh2_build(soup, num_tris, &bvh));
  printf("BVH2: triangles=%u, nodes=%u, %.1f ms\n", bvh.num_tris, bvh.num_nodes, now_ms() - t0);

  /* upload program: a pipeline descriptor stands where the .vxbin stands in the reference */
  printf("upload program\n");
  image.min_vma = 0x80000000ull;
  image.max_vma = 0x80000000ull + 64;
  image.magic = SRT_KERNEL_MAGIC;
  image.pipeline = SRT_PIPELINE_RAYTRACE;
  image.version = 1;
  RT_CHECK(vx_upload_kernel_bytes(device, &image, sizeof(image), &krnl_buffer));

  /* multi-GPU: join the communicator (collective) */
  if (num_gpus > 1) {
    unsigned char id[VX_CUDA_COMM_ID_BYTES];
    const char* id_file = getenv("SRT_COMM_ID_FILE");
    if (!id_file) { printf("Error: SRT_COMM_ID_FILE is not set\n"); cleanup(); return -1; }
    RT_CHECK(exchange_comm_id(id_file, id));
    RT_CHECK(vx_cuda_comm_init(device, id, rank_id, num_gpus, &comm));
    printf("rank %u of %u joined\n", rank_id, num_gpus);
  }

  /* allocate device memory: this rank's strip of the frame (the whole frame on one GPU) */
  strip_rows = srt_rank_rows(dst_height, ROWS_PER_BLOCK, num_gpus, rank_id);
  cbuf_size = (strip_rows ? strip_rows : 1) * dst_width * 4;
  RT_CHECK(vx_mem_alloc(device, (uint64_t)bvh.num_nodes * sizeof(srt_bvh2_node_t), VX_MEM_READ, &node_buffer));
  RT_CHECK(vx_mem_alloc(device, (uint64_t)bvh.num_tris * sizeof(srt_tri_rec_t), VX_MEM_READ, &tri_buffer));
  RT_CHECK(vx_mem_alloc(device, cbuf_size, VX_MEM_READ_WRITE, &color_buffer));
  frame_size = dst_height * dst_width * 4;
  if (num_gpus > 1 && rank_id == 0) RT_CHECK(vx_mem_alloc(device, frame_size, VX_MEM_READ_WRITE, &frame_buffer));
  RT_CHECK(vx_copy_to_dev(node_buffer, bvh.nodes, 0, (uint64_t)bvh.num_nodes * sizeof(srt_bvh2_node_t)));
  RT_CHECK(vx_copy_to_dev(tri_buffer, bvh.tris, 0, (uint64_t)bvh.num_tris * sizeof(srt_tri_rec_t)));

  /* clear destination buffer (draw3d/main.cpp:483-490) */
  printf("clear destination buffer\n");
  host_fb = (uint32_t*)malloc(frame_size > cbuf_size ? frame_size : cbuf_size);
  if (!host_fb) { cleanup(); return -1; }
  { uint32_t i; for (i = 0; i < cbuf_size / 4; ++i) host_fb[i] = 0xff000000u; }
  RT_CHECK(vx_copy_to_dev(color_buffer, host_fb, 0, cbuf_size));

  /* kernel arguments */
  kernel_arg.num_tasks = dst_height;
  kernel_arg.dst_width = dst_width;
  kernel_arg.dst_height = dst_height;
  kernel_arg.cbuf_stride = 4;
  kernel_arg.cbuf_pitch = dst_width * 4;
  kernel_arg.mode = mode;
  kernel_arg.spp = spp;
  kernel_arg.seed = seed;
  kernel_arg.num_nodes = bvh.num_nodes;
  kernel_arg.num_tris = bvh.num_tris;
  kernel_arg.rank_id = rank_id;
  kernel_arg.num_ranks = num_gpus;
  kernel_arg.rows_per_block = num_gpus > 1 ? ROWS_PER_BLOCK : 1;
  RT_CHECK(vx_mem_address(node_buffer, &addr)); kernel_arg.node_addr = addr;
  RT_CHECK(vx_mem_address(tri_buffer, &addr)); kernel_arg.tri_addr = addr;
  RT_CHECK(vx_mem_address(color_buffer, &addr)); kernel_arg.cbuf_addr = addr;
  printf("color_buffer=0x%llx\n", (unsigned long long)addr);
  RT_CHECK(vx_upload_bytes(device, &kernel_arg, sizeof(kernel_arg), &args_buffer));

  /* first run includes the one-time device-layout build; the second is the steady-state frame */
  {
    int pass;
    for (pass = 0; pass < 2; ++pass) {
      printf("start device\n");
      t0 = now_ms();
      RT_CHECK(vx_start(device, krnl_buffer, args_buffer));
      if (num_gpus > 1) /* strips -> rank 0, queued behind the frame on the launch stream */
        RT_CHECK(vx_cuda_frame_gather(device, comm, color_buffer, frame_buffer, dst_width, dst_height, ROWS_PER_BLOCK, 0));
      printf("wait for completion\n");
      RT_CHECK(vx_ready_wait(device, VX_MAX_TIMEOUT));
      t1 = now_ms();
      printf("Elapsed time: %.3f ms%s\n", t1 - t0, pass == 0 ? " (first run: includes scene repack)" : "");
    }
  }
  RT_CHECK(vx_mpm_query(device, VX_CSR_MINSTRET, 0, &v));
  printf("rays traced%s: %llu\n", num_gpus > 1 ? " by this rank" : "", (unsigned long long)v);
  RT_CHECK(vx_mpm_query(device, VX_CSR_RT_ELAPSED_NS, 0, &v));
  printf("device time: %.3f ms\n", (double)v * 1e-6);

  /* download destination buffer and save it bottom-up (draw3d/main.cpp:380-387) */
  if (rank_id != 0) { /* rank 0 holds the frame */
    cleanup();
    return 0;
  }
  printf("download destination buffer\n");
  RT_CHECK(vx_copy_from_dev(host_fb, num_gpus > 1 ? frame_buffer : color_buffer, 0, frame_size));
  if (png_save_argb(output_file, (const uint8_t*)host_fb + (int64_t)(dst_height - 1) * kernel_arg.cbuf_pitch,
                    dst_width, dst_height, -(int64_t)kernel_arg.cbuf_pitch) != 0) {
    printf("Error: cannot write %s\n", output_file);
    cleanup();
    return -1;
  }

  printf("cleanup\n");
  cleanup();

  if (reference_file) {
    int errors = png_compare(output_file, reference_file, 0);
    if (0 == errors) {
      printf("PASSED!\n");
    } else {
      printf("FAILED! %d errors.\n", errors);
      return errors < 0 ? 1 : errors;
    }
  }
  return 0;
}
