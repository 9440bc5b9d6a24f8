/* ahk_cpu.h — the CPU twin of the engine's C ABI ("identical ahk_cpu_* entry points for the CPU baseline").
 *
 * Same structs (ahk_config, ahk_buffers, ahk_step_log), same buffer layouts and the same call protocol as include/ahk.h,
 * with HOST pointers and the `stream` argument ignored.  Implemented in oracle/ahk_oracle.c over the CPU restatement of
 * the path (per env: air_hockey_env.py:26-155 + processor.py:31-66 over the SPEC.md simulator), stepped by a persistent
 * pthread pool; built into oracle/libahk_oracle.so.  It is the checker and the timed CPU baseline (bench.py --impl
 * reference), NOT part of the product: lib/libahk.so neither links nor loads it, and has no CPU fallback.
 *
 * Limits: the 3-slot ring only (ring_depth 0 or 3); no drawlist / warm buffers. */
#ifndef AHK_CPU_H
#define AHK_CPU_H

#include "ahk.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ahk_cpu_env ahk_cpu_env;

/* ≙ ahk_create / ahk_destroy (AirHockeyEnv.__init__, air_hockey_env.py:9-24).  cfg->device is ignored.
 * Worker threads: every online core, or the environment variable AHK_CPU_THREADS. */
int ahk_cpu_create(const ahk_config *cfg, ahk_cpu_env **out);
void ahk_cpu_destroy(ahk_cpu_env *env);
int ahk_cpu_threads(const ahk_cpu_env *env);

/* ≙ ahk_set_table_image: host uint8[700][450][3]; the CPU path draws the native 450x700 frame and resizes it with
 * NEAREST (processor.py:34-35), as the reference does. */
int ahk_cpu_set_table_image(ahk_cpu_env *env, const uint8_t *rgb_host, int height, int width);

/* ≙ ahk_bind_buffers, host memory: state [16][n], frames [3][n][3][128][128], reward, done, robot_label, human_label
 * (required), info, stats (optional). */
int ahk_cpu_bind_buffers(ahk_cpu_env *env, const ahk_buffers *bufs_host);

/* ≙ ahk_reset (AirHockeyEnv.reset, air_hockey_env.py:26-76) */
int ahk_cpu_reset(ahk_cpu_env *env, const uint8_t *mask, const uint8_t *episode, const uint8_t *use, void *stream_ignored);

/* ≙ ahk_step (AirHockeyEnv.step, air_hockey_env.py:84-155) */
int ahk_cpu_step(ahk_cpu_env *env, const uint8_t *robot, const uint8_t *human, void *stream_ignored);

/* ≙ ahk_step_many (the generate_data.py:44-58 loop) */
int ahk_cpu_step_many(ahk_cpu_env *env, int steps, const ahk_step_log *log_host, void *stream_ignored);

/* ≙ ahk_observe (DataProcessor.process_observation's output formats, processor.py:39-64), host memory */
int ahk_cpu_observe(ahk_cpu_env *env, int format, void *out_host, void *stream_ignored);

/* The reference's env.step() returns the processed observation every step (air_hockey_env.py:153-155).  With a host buffer bound,
 * every ahk_cpu_step / ahk_cpu_step_many also writes each env's observation in `format` (the bytes ahk_cpu_observe would write),
 * by the worker thread that just stepped the env — the reference's per-step work in one pass.  NULL unbinds.  No GPU counterpart
 * (there the frames stay uint8 in the ring and ahk_observe materialises them on request). */
int ahk_cpu_bind_observation(ahk_cpu_env *env, int format, void *out_host);

/* ≙ ahk_head */
int ahk_cpu_head(const ahk_cpu_env *env);

const char *ahk_cpu_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
