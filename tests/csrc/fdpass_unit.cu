// CPU unit test of the descriptor hand-over that symmetric windows use (csrc/p2p_fdpass.cu): a child process
// serves two descriptors (memfds with known contents) the way an exporting rank serves its VMM handles; the parent
// fetches both by name, the way an importing rank does, and reads them.  Also: a fetch from a name nobody serves
// fails after its timeout instead of hanging.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>
#include "p2p_internal.h"

static int memfd_with(const char* text) {
    int fd = memfd_create("fdpass_unit", 0);
    if (fd < 0 || write(fd, text, strlen(text)) != (ssize_t)strlen(text)) { perror("memfd"); exit(2); }
    return fd;
}

int main() {
    int pipefd[2];
    if (pipe(pipefd) != 0) return 2;
    pid_t child = fork();
    if (child == 0) {
        close(pipefd[0]);
        const int fds[2] = {memfd_with("payload of descriptor zero"), memfd_with("multicast object stand-in")};
        char name[40];
        p2p::FdServer* srv = p2p::fdpass_serve(fds, 2, 3, 10000, name);
        if (!srv) _exit(3);
        if (write(pipefd[1], name, sizeof name) != (ssize_t)sizeof name) _exit(4);
        p2p::fdpass_finish(srv);  // returns once the three requests have been served
        _exit(0);
    }
    close(pipefd[1]);
    char name[40] = {};
    if (read(pipefd[0], name, sizeof name) != (ssize_t)sizeof name) { printf("no name from the child\n"); return 1; }
    const char* want[3] = {"payload of descriptor zero", "multicast object stand-in", "payload of descriptor zero"};
    const int which[3] = {0, 1, 0};
    for (int i = 0; i < 3; ++i) {
        const int fd = p2p::fdpass_fetch(name, which[i], 5000);
        if (fd < 0) { printf("fetch %d failed\n", i); return 1; }
        char buf[64] = {};
        if (pread(fd, buf, sizeof buf - 1, 0) <= 0 || strcmp(buf, want[i]) != 0) { printf("descriptor %d reads '%s'\n", i, buf); return 1; }
        close(fd);
    }
    int status = 0;
    waitpid(child, &status, 0);
    if (!WIFEXITED(status) || WEXITSTATUS(status) != 0) { printf("server exited with %d\n", status); return 1; }
    if (p2p::fdpass_fetch("b200p2p-fd-nobody-serves-this", 0, 300) >= 0) { printf("fetch from nowhere succeeded\n"); return 1; }
    if (p2p::fdpass_fetch(name, 7, 300) >= 0) { printf("fetch after the server ended succeeded\n"); return 1; }
    printf("fdpass ok\n");
    return 0;
}
