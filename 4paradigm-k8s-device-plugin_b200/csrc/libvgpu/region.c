/* region.c — shared-cache file: create/attach, lock protocol, process slots, usage counters.
 * Mirrors the reference's protocol operation by operation (citations inline) so that the
 * resulting file is byte-identical for identical traces (tests/test_quota_parity.py). */
#define _GNU_SOURCE
#include "region.h"

#include <errno.h>
#include <fcntl.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include "shimlog.h"

static vgpu_region_t *g_reg;
static int g_fd = -1;
static int g_pid;

vgpu_region_t *region_get(void) { return g_reg; }

/* proc_alive@0x41618: reads /proc/<pid>/stat; state Z, X or x — or no such file — is dead */
static int proc_is_dead(int pid) {
    char path[64], comm[4096], state = 0;
    int p = 0;
    snprintf(path, sizeof(path), "/proc/%d/stat", pid);
    FILE *f = fopen(path, "r");
    if (!f) return 1;
    int n = fscanf(f, "%d %4095s %c", &p, comm, &state);
    fclose(f);
    if (n != 3) return 0; /* the reference returns 2 ("unknown") here and only kicks on 1 */
    return state == 'Z' || state == 'X' || state == 'x';
}

/* lock_shrreg@0x4426a: sem_timedwait 10 s, up to 30 tries; a dead owner is evicted under
 * lockf (fix_lock_shrreg@0x43c13). */
void region_lock(void) {
    vgpu_region_t *r = g_reg;
    for (int tries = 0; tries < 30; ++tries) {
        struct timespec ts;
        clock_gettime(CLOCK_REALTIME, &ts);
        ts.tv_sec += 10;
        if (sem_timedwait(&r->sem, &ts) == 0) {
            r->owner_pid = (uint64_t)g_pid;
            __sync_synchronize();
            return;
        }
        if (errno == EINTR) { --tries; continue; }
        if (errno != ETIMEDOUT) { SHIM_ERR("lock shrreg: errno=%d", errno); continue; }
        int owner = (int)r->owner_pid;
        if (owner != 0 && proc_is_dead(owner)) {
            SHIM_WARN("owner %d of the shrreg lock is dead, taking the lock over", owner);
            if (lockf(g_fd, F_LOCK, VGPU_REGION_SIZE) == 0) {
                if ((int)r->owner_pid == owner) {  /* nobody fixed it meanwhile */
                    r->owner_pid = (uint64_t)g_pid;
                    lockf(g_fd, F_ULOCK, VGPU_REGION_SIZE);
                    return;
                }
                lockf(g_fd, F_ULOCK, VGPU_REGION_SIZE);
            }
        }
    }
    SHIM_ERR("fail to lock shrreg in %d seconds", 300);
    exit(-1);
}

void region_unlock(void) {  /* unlock_shrreg@0x44652 */
    __sync_synchronize();
    g_reg->owner_pid = 0;
    sem_post(&g_reg->sem);
}

/* clear_proc_slot_nolock@0x4468b: kick slots whose pid is dead: the last slot is copied over
 * the dead one (its own bytes stay behind, stale, as in the reference). */
static int kick_dead_nolock(int self) {
    vgpu_region_t *r = g_reg;
    int i = 0, kicked = 0;
    while (i < r->proc_num) {
        int pid = r->procs[i].pid;
        if (pid != 0 && pid != self && proc_is_dead(pid)) {
            SHIM_WARN("Kick dead proc %d", pid);
            r->proc_num -= 1;
            memcpy(&r->procs[i], &r->procs[r->proc_num], sizeof(vgpu_proc_slot_t));
            __sync_synchronize();
            ++kicked;
            continue; /* re-examine the slot that was moved in */
        }
        ++i;
    }
    return kicked;
}

/* rm_quitted_process@0x425b6, the OOM path's sweep.  Not the same compaction as the one at start-up: survivors keep
 * their order and are moved down slot by slot, and only `pid` and `used[]` travel (memcpy of 0x280 bytes @0x42900) —
 * hostpid, monitorused and status of the slot they land in stay what they were; slots past the new proc_num keep
 * their bytes.  The reference decides "alive" by finding the pid in the output of `ps ax`; here /proc/<pid>/stat. */
int region_reap_dead(void) {
    region_lock();
    vgpu_region_t *r = g_reg;
    int j = 0, any = 0;
    for (int i = 0; i < r->proc_num; ++i) {
        const int pid = r->procs[i].pid;
        if (pid == 0 || (pid != g_pid && proc_is_dead(pid))) {
            SHIM_WARN("Kick dead proc %d", pid);
            any = 1;
            continue;
        }
        if (j != i) {
            r->procs[j].pid = pid;
            memcpy(r->procs[j].used, r->procs[i].used, sizeof(r->procs[j].used));
        }
        ++j;
    }
    r->proc_num = j;
    __sync_synchronize();
    region_unlock();
    return any;
}

void (*region_status_hook)(int status);  /* set by the shim; must be async-signal-safe */

static void set_my_status(int st) {
    if (g_reg)
        for (int i = 0; i < g_reg->proc_num; ++i)
            if (g_reg->procs[i].pid == g_pid) g_reg->procs[i].status = st;
    if (region_status_hook) region_status_hook(st);
}
static void on_sigusr2(int s) { (void)s; set_my_status(2); }  /* sig_swap_stub: suspend */
static void on_sigusr1(int s) { (void)s; set_my_status(1); }  /* sig_restore_stub: resume */

/* init_proc_slot_withlock@0x44820 */
static void claim_slot(void) {
    vgpu_region_t *r = g_reg;
    region_lock();
    if (r->proc_num > VGPU_MAX_PROCS - 1) {
        region_unlock();
        SHIM_ERR("too many processes share %s", "the shrreg");
        exit(-1);
    }
    signal(SIGUSR2, on_sigusr2);  /* suspend: status 2 (sig_swap_stub) */
    signal(SIGUSR1, on_sigusr1);  /* resume:  status 1 (sig_restore_stub) */
    int found = 0;
    for (int i = 0; i < r->proc_num; ++i)
        if (r->procs[i].pid == g_pid) {
            r->procs[i].status = 1;
            memset(r->procs[i].used, 0, sizeof(r->procs[i].used));
            found = 1;
            break;
        }
    if (!found) {
        vgpu_proc_slot_t *s = &r->procs[r->proc_num];
        s->pid = g_pid;
        s->status = 1;
        memset(s->used, 0, sizeof(s->used));
        r->proc_num += 1;
    }
    kick_dead_nolock(g_pid);
    region_unlock();
}

/* exit_handler@0x43f98: 3 s timed lock; zero own `used`, move the last slot into ours. */
void region_exit_handler(void) {
    vgpu_region_t *r = g_reg;
    if (!r) return;
    struct timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    ts.tv_sec += 3;
    if (sem_timedwait(&r->sem, &ts) != 0) return;
    r->owner_pid = (uint64_t)g_pid;
    for (int i = 0; i < r->proc_num; ++i)
        if (r->procs[i].pid == g_pid) {
            memset(r->procs[i].used, 0, sizeof(r->procs[i].used));
            r->proc_num -= 1;
            memcpy(&r->procs[i], &r->procs[r->proc_num], sizeof(vgpu_proc_slot_t));
            break;
        }
    __sync_synchronize();
    r->owner_pid = 0;
    sem_post(&r->sem);
}

static void child_after_fork(void) {  /* child_reinit_flag@0x449eb: the child is a new process */
    g_pid = getpid();
    if (g_reg) claim_slot();
}

int region_attach(const char *path, const vgpu_region_init_t *init) {
    g_pid = getpid();
    umask(0);
    g_fd = open(path, O_RDWR | O_CREAT, 0666);
    if (g_fd == -1) {
        SHIM_ERR("Fail to open shrreg %s: errno=%d", path, errno);
        return -errno;
    }
    if (lseek(g_fd, VGPU_REGION_SIZE, SEEK_SET) != (off_t)VGPU_REGION_SIZE || write(g_fd, "", 1) != 1 ||
        lseek(g_fd, 0, SEEK_SET) != 0) {
        SHIM_ERR("Fail to size shrreg %s: errno=%d", path, errno);
        return -errno;
    }
    void *m = mmap(NULL, VGPU_REGION_SIZE, PROT_READ | PROT_WRITE, MAP_SHARED, g_fd, 0);
    if (m == MAP_FAILED) {
        SHIM_ERR("Fail to map shrreg %s: errno=%d", path, errno);
        return -errno;
    }
    g_reg = (vgpu_region_t *)m;
    if (lockf(g_fd, F_LOCK, VGPU_REGION_SIZE) != 0) {
        SHIM_ERR("Fail to lock shrreg %s: errno=%d", path, errno);
        return -errno;
    }
    vgpu_region_t *r = g_reg;
    /* put_device_info@0x41e2e runs before the magic check, on every attach */
    r->device_num = init->device_num;
    for (uint64_t i = 0; i < init->device_num && i < VGPU_MAX_DEVICES; ++i)  /* NVML writes the string only: bytes after its terminator stay */
        memcpy(r->uuids[i], init->uuids[i], strnlen(init->uuids[i], VGPU_UUID_LEN - 1) + 1);
    if (r->initialized_flag != VGPU_REGION_MAGIC) {
        memcpy(r->limit, init->limit, sizeof(r->limit));
        memcpy(r->sm_limit, init->sm_limit, sizeof(r->sm_limit));
        if (sem_init(&r->sem, 1, 1) != 0) SHIM_ERR("Fail to init sem %s: errno=%d", path, errno);
        r->sm_init_flag = 0;
        r->utilization_switch = 1;
        r->recent_kernel = 2;
        r->priority = init->priority;
        __sync_synchronize();
        r->initialized_flag = VGPU_REGION_MAGIC;
    } else {
        for (int i = 0; i < VGPU_MAX_DEVICES; ++i) {
            if (r->limit[i] != init->limit[i])
                SHIM_ERR("Limit inconsistency detected for %dth device, %lu expected, get %lu", i,
                         (unsigned long)r->limit[i], (unsigned long)init->limit[i]);
            if (r->sm_limit[i] != init->sm_limit[i])
                SHIM_INFO("SM limit inconsistency detected for %dth device, %lu expected, get %lu", i,
                          (unsigned long)r->sm_limit[i], (unsigned long)init->sm_limit[i]);
        }
    }
    if (lockf(g_fd, F_ULOCK, VGPU_REGION_SIZE) != 0) SHIM_ERR("Fail to unlock shrreg %s: errno=%d", path, errno);
    claim_slot();
    atexit(region_exit_handler);
    pthread_atfork(NULL, NULL, child_after_fork);
    return 0;
}

static vgpu_proc_slot_t *my_slot(void) {
    vgpu_region_t *r = g_reg;
    for (int i = 0; i < r->proc_num; ++i)
        if (r->procs[i].pid == g_pid) return &r->procs[i];
    return NULL;
}

int region_add_usage(int dev, uint64_t bytes, int kind) {
    if (dev < 0 || dev >= VGPU_MAX_DEVICES) return -1;
    region_lock();
    vgpu_proc_slot_t *s = my_slot();
    if (s) {
        s->used[dev].total += bytes;
        if (kind == VGPU_USE_CONTEXT) s->used[dev].context_size += bytes;
        else if (kind == VGPU_USE_MODULE) s->used[dev].module_size += bytes;
        else s->used[dev].data_size += bytes;
    }
    region_unlock();
    return s ? 0 : -1;
}

int region_rm_usage(int dev, uint64_t bytes, int kind) {
    if (dev < 0 || dev >= VGPU_MAX_DEVICES) return -1;
    region_lock();
    vgpu_proc_slot_t *s = my_slot();
    if (s) {
        s->used[dev].total -= bytes;
        if (kind == VGPU_USE_CONTEXT) s->used[dev].context_size -= bytes;
        else if (kind == VGPU_USE_MODULE) s->used[dev].module_size -= bytes;
        else s->used[dev].data_size -= bytes;
    }
    region_unlock();
    return s ? 0 : -1;
}

uint64_t region_usage(int dev) {
    if (dev < 0 || dev >= VGPU_MAX_DEVICES) return 0;
    region_lock();
    uint64_t t = 0;
    for (int i = 0; i < g_reg->proc_num; ++i) t += g_reg->procs[i].used[dev].total;
    region_unlock();
    return t;
}

uint64_t region_limit(int dev) {
    /* get_current_device_memory_limit@0x45ae1 rejects dev > 16, not >= 16: for dev == 16 it reads one element past
     * limit[], which is sm_limit[0] in the file layout.  Same answer here, without the out-of-bounds access. */
    if (dev == VGPU_MAX_DEVICES) return g_reg->sm_limit[0];
    if (dev < 0 || dev > VGPU_MAX_DEVICES) { SHIM_ERR("Illegal device id: %d", dev); return 0; }
    return g_reg->limit[dev];
}
uint64_t region_sm_limit(int dev) { return (dev < 0 || dev >= VGPU_MAX_DEVICES) ? 0 : g_reg->sm_limit[dev]; }

void region_set_hostpid(int hostpid) {
    if (!g_reg) return;
    region_lock();
    vgpu_proc_slot_t *s = my_slot();
    if (s) {
        s->hostpid = hostpid;
        /* set_host_pid@0x455d8 also wipes what a monitor had reported for whoever held this slot before (loop @0x456d1) */
        memset(s->monitorused, 0, sizeof(s->monitorused));
    }
    region_unlock();
}

void region_set_my_status(int status) {
    if (!g_reg) return;
    region_lock();
    vgpu_proc_slot_t *s = my_slot();
    if (s) s->status = status;
    region_unlock();
}

int region_my_status(void) {
    vgpu_proc_slot_t *s = g_reg ? my_slot() : NULL;
    return s ? s->status : 1;
}
