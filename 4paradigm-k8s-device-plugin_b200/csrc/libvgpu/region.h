/*
 * region.h — the cross-process shared-cache file (Contract C).
 *
 * Byte-for-byte the layout the reference writes (verified against a file
 * the reference produced: libvgpu.so try_create_shrreg@0x44bcf), so that old and new
 * libraries — and any external monitor — can share one CUDA_DEVICE_MEMORY_SHARED_CACHE file.
 */
#ifndef VGPU_REGION_H_
#define VGPU_REGION_H_
#include <semaphore.h>
#include <stdint.h>

#define VGPU_MAX_DEVICES 16          /* "Max Gpus Per Node can't excced 16", @0x27f0d */
#define VGPU_MAX_PROCS   1024
#define VGPU_REGION_MAGIC 19920718   /* 0x12FF74E, @0x450a5 */
#define VGPU_REGION_SIZE 0xC4748u    /* + 1 trailing byte written at lseek(SIZE) */
#define VGPU_UUID_LEN 96

typedef struct {
    uint64_t context_size, module_size, data_size, offset, total;
} vgpu_dev_usage_t;                              /* 0x28 */

typedef struct {
    int32_t pid;                                 /* +0x000 */
    int32_t hostpid;                             /* +0x004 */
    vgpu_dev_usage_t used[VGPU_MAX_DEVICES];     /* +0x008, 0x280 bytes */
    uint64_t monitorused[VGPU_MAX_DEVICES];      /* +0x288 */
    int32_t status;                              /* +0x308: 1 running, 2 suspended */
    int32_t pad;
} vgpu_proc_slot_t;                              /* 0x310 */

typedef struct {
    int32_t initialized_flag;                    /* +0x00 */
    int32_t sm_init_flag;                        /* +0x04 */
    uint64_t owner_pid;                          /* +0x08 */
    sem_t sem;                                   /* +0x10, 32 bytes, pshared */
    uint64_t device_num;                         /* +0x30 */
    char uuids[VGPU_MAX_DEVICES][VGPU_UUID_LEN]; /* +0x38 */
    uint64_t limit[VGPU_MAX_DEVICES];            /* +0x638 */
    uint64_t sm_limit[VGPU_MAX_DEVICES];         /* +0x6B8 */
    vgpu_proc_slot_t procs[VGPU_MAX_PROCS];      /* +0x738 */
    int32_t proc_num;                            /* +0xC4738 */
    int32_t utilization_switch;                  /* +0xC473C */
    int32_t recent_kernel;                       /* +0xC4740 */
    int32_t priority;                            /* +0xC4744 */
} vgpu_region_t;

_Static_assert(sizeof(sem_t) == 32, "sem_t must be 32 bytes as in the reference build");
_Static_assert(sizeof(vgpu_proc_slot_t) == 0x310, "slot stride");
_Static_assert(__builtin_offsetof(vgpu_region_t, limit) == 0x638, "limit offset");
_Static_assert(__builtin_offsetof(vgpu_region_t, procs) == 0x738, "procs offset");
_Static_assert(__builtin_offsetof(vgpu_region_t, proc_num) == 0xC4738, "proc_num offset");
_Static_assert(sizeof(vgpu_region_t) == VGPU_REGION_SIZE, "region size");

/* kinds for add/rm usage (reference passes 0 context, 1 module, 2 data) */
enum { VGPU_USE_CONTEXT = 0, VGPU_USE_MODULE = 1, VGPU_USE_DATA = 2 };

typedef struct {
    uint64_t limit[VGPU_MAX_DEVICES];
    uint64_t sm_limit[VGPU_MAX_DEVICES];
    uint64_t device_num;
    char uuids[VGPU_MAX_DEVICES][VGPU_UUID_LEN];
    int32_t priority;
} vgpu_region_init_t;

/* Opens (creating if needed) and maps the file, initialises it if this process is the
 * first comer, claims a process slot.  Returns 0 or -errno. */
int  region_attach(const char *path, const vgpu_region_init_t *init);
vgpu_region_t *region_get(void);
void region_lock(void);
void region_unlock(void);
void region_exit_handler(void);                  /* exit_handler@0x43f98 */
int  region_add_usage(int dev, uint64_t bytes, int kind);   /* add_gpu_device_memory_usage@0x43536 */
int  region_rm_usage(int dev, uint64_t bytes, int kind);    /* rm_gpu_device_memory_usage@0x43909 */
uint64_t region_usage(int dev);                  /* get_gpu_memory_usage@0x42be5 */
uint64_t region_limit(int dev);                  /* get_current_device_memory_limit@0x45ae1 */
uint64_t region_sm_limit(int dev);
int  region_reap_dead(void);                     /* rm_quitted_process@0x425b6; returns #reaped */
int  region_my_status(void);
void region_set_my_status(int status);           /* set_current_gpu_status@0x45c9d */
void region_set_hostpid(int hostpid);            /* set_host_pid@0x4497c */
extern void (*region_status_hook)(int status);  /* called from the SIGUSR1/2 handlers */

#endif
