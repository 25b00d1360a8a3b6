/* c_driver.c -- a compiled caller of the C ABI (include/brax_b200.h) with no Python and no
 * torch in the process: what a cgo / JNI / XLA-FFI binding does, reduced to plain C.
 *
 *   c_driver <system.blob> <io.blob> <out.blob>
 *
 * system.blob: the bytes of a bx_system_desc followed by its bodies / joints / actuators /
 *              forces / pairs arrays and, if desc.env / desc.reset were set, one bx_env_desc / one
 *              bx_reset_desc (pointer fields are patched here) -- tests/test_cdriver.py writes it from a SystemDesc.
 * io.blob:     int64 B, int64 T, then pos[B,N,3] rot[B,N,4] vel[B,N,3] ang[B,N,3] act[T,B,A].
 * out.blob:    after T chained bx_step_host calls (System.step, system.py:244, fed with its own
 *              QP): pos rot vel ang of the last state, then the Info of the last step
 *              (contact.vel/.ang, joint.vel/.ang, actuator.vel/.ang [B,N,3] each, contact_pos
 *              [B,K,3], contact_normal [B,K,3], contact_penetration [B,K]) and, with an env
 *              epilogue, obs [B,O] reward [B] done [B] metrics [B,M] of the last step.
 * Buffers are page-locked (bx_host_alloc): the step works on them in place over PCIe. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "brax_b200.h"

static void die(const char* what) {
  fprintf(stderr, "c_driver: %s: %s\n", what, bx_last_error());
  exit(1);
}

static void* slurp(FILE* f, size_t bytes) {
  void* p = malloc(bytes ? bytes : 1);
  if (bytes && fread(p, 1, bytes, f) != bytes) { fprintf(stderr, "c_driver: short read\n"); exit(1); }
  return p;
}

static float* pinned(size_t floats) {
  float* p = (float*)bx_host_alloc(floats * sizeof(float));
  if (!p) die("bx_host_alloc");
  memset(p, 0, floats * sizeof(float));
  return p;
}

int main(int argc, char** argv) {
  if (argc != 4) { fprintf(stderr, "usage: c_driver system.blob io.blob out.blob\n"); return 2; }
  FILE* f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 1; }
  bx_system_desc* d = (bx_system_desc*)slurp(f, sizeof(bx_system_desc));
  const int has_env = d->env != NULL, has_reset = d->reset != NULL;
  d->bodies = (const bx_body*)slurp(f, sizeof(bx_body) * (size_t)d->num_bodies);
  d->joints = (const bx_joint*)slurp(f, sizeof(bx_joint) * (size_t)d->num_joints);
  d->actuators = (const bx_actuator*)slurp(f, sizeof(bx_actuator) * (size_t)d->num_actuators);
  d->forces = (const bx_force*)slurp(f, sizeof(bx_force) * (size_t)d->num_forces);
  d->pairs = (const bx_pair*)slurp(f, sizeof(bx_pair) * (size_t)d->num_pairs);
  d->env = has_env ? (const bx_env_desc*)slurp(f, sizeof(bx_env_desc)) : NULL;
  d->reset = has_reset ? (const bx_reset_desc*)slurp(f, sizeof(bx_reset_desc)) : NULL;
  fclose(f);

  bx_system* sys = NULL;
  if (bx_system_create(d, 0, &sys)) die("bx_system_create");
  const size_t n = (size_t)bx_system_num_bodies(sys), a = (size_t)bx_system_act_size(sys);
  const size_t k = (size_t)bx_system_num_contacts(sys);
  const size_t o = (size_t)bx_system_obs_size(sys), m = (size_t)bx_system_num_metrics(sys);

  f = fopen(argv[2], "rb");
  if (!f) { perror(argv[2]); return 1; }
  int64_t hdr[2];
  if (fread(hdr, sizeof(int64_t), 2, f) != 2) { fprintf(stderr, "c_driver: short read\n"); return 1; }
  const size_t b = (size_t)hdr[0], t_n = (size_t)hdr[1];
  /* two QP sets: the step reads one and writes the other */
  float* qp[2][4];
  const size_t w[4] = {3, 4, 3, 3};
  for (int s = 0; s < 2; s++)
    for (int i = 0; i < 4; i++) qp[s][i] = pinned(b * n * w[i]);
  float* act = pinned(t_n * b * a + 1);
  for (int i = 0; i < 4; i++)
    if (fread(qp[0][i], sizeof(float), b * n * w[i], f) != b * n * w[i]) return 1;
  if (fread(act, sizeof(float), t_n * b * a, f) != t_n * b * a) return 1;
  fclose(f);

  float* info[9];
  const size_t iw[9] = {3 * n, 3 * n, 3 * n, 3 * n, 3 * n, 3 * n, 3 * k, 3 * k, k};
  for (int i = 0; i < 9; i++) info[i] = pinned(b * iw[i] + 1);
  float *obs = pinned(b * o + 1), *reward = pinned(b + 1), *done = pinned(b + 1), *metrics = pinned(b * m + 1);

  int cur = 0;
  for (size_t t = 0; t < t_n; t++, cur ^= 1) {
    bx_step_io io;
    memset(&io, 0, sizeof(io));
    io.pos = qp[cur][0]; io.rot = qp[cur][1]; io.vel = qp[cur][2]; io.ang = qp[cur][3];
    io.act = act + t * b * a;
    io.pos_o = qp[cur ^ 1][0]; io.rot_o = qp[cur ^ 1][1]; io.vel_o = qp[cur ^ 1][2]; io.ang_o = qp[cur ^ 1][3];
    io.contact_vel = info[0]; io.contact_ang = info[1]; io.joint_vel = info[2]; io.joint_ang = info[3];
    io.actuator_vel = info[4]; io.actuator_ang = info[5];
    io.contact_pos = info[6]; io.contact_normal = info[7]; io.contact_penetration = info[8];
    if (has_env) { io.obs = obs; io.reward = reward; io.done = done; io.metrics = metrics; }
    if (bx_step_host(sys, (int64_t)b, &io)) die("bx_step_host");
  }

  f = fopen(argv[3], "wb");
  if (!f) { perror(argv[3]); return 1; }
  for (int i = 0; i < 4; i++) fwrite(qp[cur][i], sizeof(float), b * n * w[i], f);
  for (int i = 0; i < 9; i++) fwrite(info[i], sizeof(float), b * iw[i], f);
  if (has_env) {
    fwrite(obs, sizeof(float), b * o, f); fwrite(reward, sizeof(float), b, f);
    fwrite(done, sizeof(float), b, f); fwrite(metrics, sizeof(float), b * m, f);
  }
  fclose(f);
  printf("c_driver: %zu envs x %zu steps, N=%zu A=%zu K=%zu O=%zu M=%zu, %lld launches\n", b, t_n, n, a, k, o, m,
         (long long)bx_launch_count());
  bx_system_destroy(sys);
  return 0;
}
