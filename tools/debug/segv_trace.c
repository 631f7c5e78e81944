// LD_PRELOAD helper for GPU-box debugging: prints a native backtrace on SIGSEGV (there is no gdb on the box).
// build: gcc -shared -fPIC -O1 -o gpurun_out/segv_trace.so tools/debug/segv_trace.c
#define _GNU_SOURCE
#include <execinfo.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <fcntl.h>
#include <unistd.h>
static int g_fd = 2;   /* pytest captures fd 2: SEGV_TRACE_FILE names a file of our own */
static void on_segv(int sig, siginfo_t* si, void* ctx) {
    (void)ctx;
    void* frames[64];
    int n = backtrace(frames, 64);
    dprintf(g_fd, "\n=== signal %d at address %p: native backtrace ===\n", sig, si->si_addr);
    backtrace_symbols_fd(frames, n, g_fd);
    _exit(139);
}
__attribute__((constructor)) static void install(void) {
    const char* f = getenv("SEGV_TRACE_FILE");
    if (f) { int fd = open(f, O_WRONLY | O_CREAT | O_APPEND, 0644); if (fd >= 0) g_fd = fd; }
    static char altstack[1 << 16];            /* a stack overflow leaves no room for the handler on the thread's own stack */
    stack_t ss;
    ss.ss_sp = altstack; ss.ss_size = sizeof(altstack); ss.ss_flags = 0;
    sigaltstack(&ss, NULL);
    struct sigaction sa;
    sa.sa_sigaction = on_segv;
    sigemptyset(&sa.sa_mask);
    sa.sa_flags = SA_SIGINFO | SA_RESETHAND | SA_ONSTACK;
    sigaction(SIGSEGV, &sa, NULL);
    sigaction(SIGBUS, &sa, NULL);
}
