// Passing file descriptors between the ranks' processes over Unix-domain sockets (SCM_RIGHTS).
//
// cuMemExportToShareableHandle(CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) yields a descriptor that is only valid in
// the exporting process; MPI_Allgather (or any byte transport) cannot carry it.  Each exporting rank therefore
// listens on an abstract-namespace socket whose name travels in its blob, and hands a duplicate of the descriptor
// to every peer that asks.  The server is a helper thread that lives from export to import and then ends: the data
// path has no thread of its own.  Plain host code (no CUDA), unit-tested on CPU by tests/csrc/fdpass_unit.cu.
#include "p2p_internal.h"

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

namespace p2p {

struct FdServer {
    int listen_fd = -1;
    std::vector<int> fds;
    int requests = 0;
    int timeout_ms = 30000;
    std::thread thread;
    std::atomic<int> served{0};
};

namespace {

socklen_t abstract_address(const char* name, sockaddr_un* addr) {
    memset(addr, 0, sizeof *addr);
    addr->sun_family = AF_UNIX;
    const size_t n = strnlen(name, sizeof addr->sun_path - 2);
    memcpy(addr->sun_path + 1, name, n);  // leading NUL: abstract namespace, nothing to unlink
    return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

bool send_fd(int sock, int fd) {
    char byte = 'F';
    iovec iov = {&byte, 1};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))] = {};
    msghdr msg = {};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof ctrl;
    cmsghdr* c = CMSG_FIRSTHDR(&msg);
    c->cmsg_level = SOL_SOCKET;
    c->cmsg_type = SCM_RIGHTS;
    c->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(c), &fd, sizeof(int));
    return sendmsg(sock, &msg, MSG_NOSIGNAL) == 1;
}

int recv_fd(int sock) {
    char byte = 0;
    iovec iov = {&byte, 1};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))] = {};
    msghdr msg = {};
    msg.msg_iov = &iov;
    msg.msg_iovlen = 1;
    msg.msg_control = ctrl;
    msg.msg_controllen = sizeof ctrl;
    if (recvmsg(sock, &msg, 0) != 1 || byte != 'F') return -1;
    for (cmsghdr* c = CMSG_FIRSTHDR(&msg); c; c = CMSG_NXTHDR(&msg, c))
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) {
            int fd = -1;
            memcpy(&fd, CMSG_DATA(c), sizeof(int));
            return fd;
        }
    return -1;
}

void serve(FdServer* s) {
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(s->timeout_ms);
    while (s->served.load() < s->requests) {
        const auto left = std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now()).count();
        if (left <= 0) break;
        pollfd p = {s->listen_fd, POLLIN, 0};
        if (poll(&p, 1, (int)std::min<long long>(left, 200)) <= 0) continue;
        const int c = accept(s->listen_fd, nullptr, nullptr);
        if (c < 0) continue;
        unsigned char which = 0xff;
        pollfd pc = {c, POLLIN, 0};
        if (poll(&pc, 1, 2000) > 0 && read(c, &which, 1) == 1 && which < s->fds.size() && send_fd(c, s->fds[which])) s->served++;
        close(c);
    }
}

}  // namespace

FdServer* fdpass_serve(const int* fds, int nfds, int requests, int timeout_ms, char* name_out) {
    static std::atomic<unsigned> counter{0};
    FdServer* s = new FdServer;
    s->fds.assign(fds, fds + nfds);
    s->requests = requests;
    s->timeout_ms = timeout_ms;
    s->listen_fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (s->listen_fd < 0) { delete s; return nullptr; }
    snprintf(name_out, 40, "b200p2p-fd-%d-%u-%llx", (int)getpid(), counter++,
             (unsigned long long)std::chrono::steady_clock::now().time_since_epoch().count() & 0xffffffull);
    sockaddr_un addr;
    const socklen_t len = abstract_address(name_out, &addr);
    if (bind(s->listen_fd, (sockaddr*)&addr, len) != 0 || listen(s->listen_fd, 64) != 0) {
        close(s->listen_fd);
        delete s;
        return nullptr;
    }
    if (requests > 0) s->thread = std::thread(serve, s);
    return s;
}

int fdpass_fetch(const char* name, int which, int timeout_ms) {
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
    sockaddr_un addr;
    const socklen_t len = abstract_address(name, &addr);
    for (;;) {
        const int c = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (c < 0) return -1;
        if (connect(c, (sockaddr*)&addr, len) == 0) {
            const unsigned char w = (unsigned char)which;
            int fd = -1;
            pollfd p = {c, POLLIN, 0};
            if (write(c, &w, 1) == 1 && poll(&p, 1, timeout_ms) > 0) fd = recv_fd(c);
            close(c);
            return fd;
        }
        close(c);
        if (std::chrono::steady_clock::now() >= deadline) return -1;  // the peer's server is not up (yet)
        usleep(1000);
    }
}

void fdpass_finish(FdServer* s) {
    if (!s) return;
    if (s->thread.joinable()) s->thread.join();
    if (s->listen_fd >= 0) close(s->listen_fd);
    delete s;
}

}  // namespace p2p
