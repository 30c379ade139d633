/* shimlog.h — stderr logging in the reference's format and level scheme
 * (LIBCUDA_LOG_LEVEL, README.md:224-233; thresholds SURVEY.md §4).  The level is read once,
 * not at every log site as the reference does (getenv+atoi per macro). */
#ifndef VGPU_SHIMLOG_H_
#define VGPU_SHIMLOG_H_
#include <pthread.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

int shim_log_level(void);   /* -1 when LIBCUDA_LOG_LEVEL is unset */
#define SHIM_BASENAME (strrchr(__FILE__, '/') ? strrchr(__FILE__, '/') + 1 : __FILE__)
#define SHIM_ERR(fmt, ...)  fprintf(stderr, "[4pdvGPU ERROR (pid:%d thread=%ld %s:%d)]: " fmt "\n", getpid(), (long)pthread_self(), SHIM_BASENAME, __LINE__, ##__VA_ARGS__)
#define SHIM_LOGAT(min, tag, fmt, ...) do { if (shim_log_level() > (min)) fprintf(stderr, "[4pdvGPU " tag "(%d:%ld:%s:%d)]: " fmt "\n", getpid(), (long)pthread_self(), SHIM_BASENAME, __LINE__, ##__VA_ARGS__); } while (0)
#define SHIM_WARN(fmt, ...)  SHIM_LOGAT(1, "Warn", fmt, ##__VA_ARGS__)
#define SHIM_INFO(fmt, ...)  SHIM_LOGAT(2, "Info", fmt, ##__VA_ARGS__)
#define SHIM_DEBUG(fmt, ...) SHIM_LOGAT(3, "Debug", fmt, ##__VA_ARGS__)
#endif
