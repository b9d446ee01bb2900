/* LD_PRELOAD helper for overlay/run_reftests.sh: says where one of the reference's test programs died (SIGSEGV,
 * SIGBUS, SIGABRT) -- the GPU box has no debugger.  Prints the faulting address and pc, the top stack words (a call through a null
 * pointer leaves its caller there), and the unwinder's backtrace.  Test infrastructure only. */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <execinfo.h>
#include <pthread.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <ucontext.h>
#include <unistd.h>

static void say(const char *what, void *pc)
{
    Dl_info info;
    char line[512];
    int n;
    if( dladdr(pc, &info) && NULL != info.dli_fname )
        n = snprintf(line, sizeof(line), "  %-8s %p  %s(%s+0x%lx)\n", what, pc, info.dli_fname,
                     info.dli_sname ? info.dli_sname : "?",
                     (unsigned long)((char*)pc - (char*)(info.dli_saddr ? info.dli_saddr : info.dli_fbase)));
    else
        n = snprintf(line, sizeof(line), "  %-8s %p  (not in a loaded object)\n", what, pc);
    if( write(2, line, (size_t)n) < 0 ) {}
}

static volatile int entered = 0;
static pthread_t first;

static void on_fault(int sig, siginfo_t *si, void *ctx)
{
    if( __sync_lock_test_and_set(&entered, 1) ) {           /* not the first fault */
        if( pthread_equal(first, pthread_self()) ) _exit(128 + sig);   /* the report itself faulted: give up */
        for( ;; ) pause();                                   /* another thread: let the first one finish its report */
    }
    first = pthread_self();
    ucontext_t *uc = (ucontext_t*)ctx;
    void *frames[64];
    char head[128];
    int unwind = 1;
    int n = snprintf(head, sizeof(head), "== signal %d, fault address %p\n", sig, si->si_addr);
    fflush(stdout);
    if( write(2, head, (size_t)n) < 0 ) {}
#if defined(__x86_64__)
    void *pc = (void*)uc->uc_mcontext.gregs[REG_RIP];
    void **sp = (void**)uc->uc_mcontext.gregs[REG_RSP];
    Dl_info where;
    say("pc", pc);
    unwind = dladdr(pc, &where) && NULL != where.dli_fname;
    for( int i = 0; i < 12 && NULL != sp; i++ ) say("stack", sp[i]);        /* a call through NULL leaves the caller here */
#endif
    if( unwind ) {   /* the unwinder cannot start from a pc that is not code */
        n = backtrace(frames, 64);
        backtrace_symbols_fd(frames, n, 2);
    }
    _exit(128 + sig);   /* the shell's code for a death by this signal, without another fault on the way out */
}

__attribute__((constructor)) static void install(void)
{
    struct sigaction sa;
    memset(&sa, 0, sizeof(sa));
    sa.sa_sigaction = on_fault;
    sa.sa_flags = SA_SIGINFO | SA_NODEFER;
    sigaction(SIGSEGV, &sa, NULL);
    sigaction(SIGABRT, &sa, NULL);
    sigaction(SIGBUS, &sa, NULL);
}
