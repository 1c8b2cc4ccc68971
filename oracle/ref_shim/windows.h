/*
 * Minimal <windows.h> stand-in so the reference's single translation unit
 * (src/win32_main.cpp, compiled where it lies under /root/reference) builds
 * with g++ on Linux.  TEST INFRASTRUCTURE ONLY: nothing under oracle/ is
 * linked into, or called from, the shipped CUDA library.
 *
 * Only what win32_main.cpp touches is provided:
 *   threads      CreateThread / WaitForSingleObject / ExitThread  (main.cpp:239-246, 925-932, 977, 874, 986)
 *   barriers     _WriteBarrier / _ReadBarrier / MemoryBarrier      (main.cpp:837-838)
 *   core count   GetProcAddress("GetLogicalProcessorInformation")  (main.cpp:2129-2155)
 *   min / max    the Win32 macros                                  (main.cpp:294, 898, 906, 1691, 2162-2181)
 */
#ifndef PT_REF_SHIM_WINDOWS_H
#define PT_REF_SHIM_WINDOWS_H

#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>
#include <memory>     /* std::unique_ptr is used at main.cpp:1026 without an include */
#include <atomic>

typedef unsigned int   DWORD;
typedef DWORD         *PDWORD;
typedef void          *LPVOID;
typedef int            BOOL;
typedef uintptr_t      ULONG_PTR;
typedef void          *HMODULE;
typedef void          *FARPROC;

#define WINAPI
#define _In_
#define TEXT(s) s
#define TRUE  1
#define FALSE 0
#define INFINITE      0xFFFFFFFFu
#define WAIT_OBJECT_0 0u
#define WAIT_TIMEOUT  258u
#define ERROR_INSUFFICIENT_BUFFER 122u

typedef DWORD (*LPTHREAD_START_ROUTINE)(LPVOID);

struct pt_shim_thread {
    pthread_t tid;
    LPTHREAD_START_ROUTINE fn;
    LPVOID arg;
    std::atomic<int> done;
    int joined;
};
typedef pt_shim_thread *HANDLE;

/* timing taps read by the harness: first worker start / last worker join (ns, CLOCK_MONOTONIC) */
extern "C" {
extern volatile int64_t pt_shim_first_worker_ns;
extern volatile int64_t pt_shim_last_join_ns;
extern int pt_shim_core_count;      /* what the fake GetLogicalProcessorInformation reports */
}

static inline int64_t pt_shim_now_ns() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (int64_t)ts.tv_sec * 1000000000ll + ts.tv_nsec;
}

static thread_local pt_shim_thread *pt_shim_self = nullptr;

static void *pt_shim_trampoline(void *p) {
    pt_shim_thread *t = (pt_shim_thread *)p;
    pt_shim_self = t;
    t->fn(t->arg);
    t->done.store(1);
    return nullptr;
}

/* drop-in demo (pathtracer_ref_b200): lets the wrapper TU substitute one thread routine -- the reference's
   MasterThread -- without touching the reference's source */
typedef LPTHREAD_START_ROUTINE (*pt_shim_thread_hook_t)(LPTHREAD_START_ROUTINE);
static pt_shim_thread_hook_t pt_shim_thread_hook = nullptr;

static inline HANDLE CreateThread(void *, size_t, LPTHREAD_START_ROUTINE fn, LPVOID arg, DWORD, void *) {
    pt_shim_thread *t = new pt_shim_thread();
    if (pt_shim_thread_hook) fn = pt_shim_thread_hook(fn);
    t->fn = fn;
    t->arg = arg;
    t->done.store(0);
    t->joined = 0;
    /* a thread created from a non-main thread is a render worker (MasterThread spawns them) */
    if (pt_shim_self != nullptr && pt_shim_first_worker_ns == 0) pt_shim_first_worker_ns = pt_shim_now_ns();
    pthread_attr_t attr;
    pthread_attr_init(&attr);
    pthread_attr_setstacksize(&attr, 16u << 20);
    pthread_create(&t->tid, &attr, pt_shim_trampoline, t);
    pthread_attr_destroy(&attr);
    return t;
}

static inline DWORD WaitForSingleObject(HANDLE h, DWORD ms) {
    if (ms == INFINITE) {
        if (!h->joined) { pthread_join(h->tid, nullptr); h->joined = 1; }
        if (pt_shim_self != nullptr) pt_shim_last_join_ns = pt_shim_now_ns();
        return WAIT_OBJECT_0;
    }
    return h->done.load() ? WAIT_OBJECT_0 : WAIT_TIMEOUT;
}

[[noreturn]] static inline void ExitThread(DWORD) {
    if (pt_shim_self) pt_shim_self->done.store(1);
    pthread_exit(nullptr);
}

/* used inside comma expressions (main.cpp:837-838), so they must be expressions */
static inline void _WriteBarrier() { __asm__ __volatile__("" ::: "memory"); }
static inline void _ReadBarrier()  { __asm__ __volatile__("" ::: "memory"); }
static inline void MemoryBarrier() { __sync_synchronize(); }

/* ---- core-count discovery (two-call protocol, main.cpp:2135-2139) ---- */
typedef enum { RelationProcessorCore = 0, RelationNumaNode, RelationCache, RelationProcessorPackage } LOGICAL_PROCESSOR_RELATIONSHIP;
typedef struct {
    ULONG_PTR ProcessorMask;
    LOGICAL_PROCESSOR_RELATIONSHIP Relationship;
    unsigned long long Reserved[2];
} SYSTEM_LOGICAL_PROCESSOR_INFORMATION, *PSYSTEM_LOGICAL_PROCESSOR_INFORMATION;

static thread_local DWORD pt_shim_last_error = 0;
static inline DWORD GetLastError() { return pt_shim_last_error; }

static BOOL pt_shim_glpi(PSYSTEM_LOGICAL_PROCESSOR_INFORMATION buf, PDWORD len) {
    int cores = pt_shim_core_count > 0 ? pt_shim_core_count : (int)sysconf(_SC_NPROCESSORS_ONLN);
    if (cores < 1) cores = 1;
    DWORD need = (DWORD)(cores * sizeof(SYSTEM_LOGICAL_PROCESSOR_INFORMATION));
    if (buf == nullptr || *len < need) {
        *len = need;
        pt_shim_last_error = ERROR_INSUFFICIENT_BUFFER;
        return FALSE;
    }
    for (int i = 0; i < cores; i++) {
        memset(&buf[i], 0, sizeof(buf[i]));
        buf[i].ProcessorMask = (ULONG_PTR)1 << (i & 63);
        buf[i].Relationship = RelationProcessorCore;
    }
    *len = need;
    return TRUE;
}
static inline HMODULE GetModuleHandle(const char *) { return (HMODULE)1; }
static inline FARPROC GetProcAddress(HMODULE, const char *name) {
    if (strcmp(name, "GetLogicalProcessorInformation") == 0) return (FARPROC)pt_shim_glpi;
    return nullptr;
}

/* Win32 min/max macros -- defined LAST: ray_math.hpp (included before this header,
   main.cpp:5-6) uses std::min/std::max, and every std header is already in. */
#ifndef min
#define min(a, b) (((a) < (b)) ? (a) : (b))
#endif
#ifndef max
#define max(a, b) (((a) > (b)) ? (a) : (b))
#endif

#endif
