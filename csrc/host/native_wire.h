// Wire format between the `final --gpus N` launcher (the switchboard, which never touches CUDA) and its
// N worker processes: fixed-size messages over one SOCK_SEQPACKET Unix socket pair per worker, file
// descriptors attached with SCM_RIGHTS. Every worker sends the same sequence of rounds; the
// launcher answers a round once it has heard from all workers, so every round is also a barrier.
#pragma once
#include <stdint.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/uio.h>

enum {
    SA_WIRE_FDS = 1,     /* all-gather: every rank sends ONE descriptor, everyone gets all of them in rank order */
    SA_WIRE_ROOT_FD = 2, /* broadcast: rank 0 sends zero or one descriptor, everyone gets it                       */
    SA_WIRE_MIN = 3,     /* all-reduce(min) of `value` (votes, barriers)                                            */
    SA_WIRE_DONE = 4     /* the worker's last message: `value` is its exit status; not answered                    */
};
#define SA_WIRE_MAX_FDS 16

typedef struct sa_wire_msg {
    uint32_t type, rank;
    int64_t value;
    uint32_t nfds, pad;
} sa_wire_msg;

static inline int sa_wire_send(int sock, const sa_wire_msg* m, const int* fds) {
    struct iovec iov;
    iov.iov_base = (void*)m;
    iov.iov_len = sizeof(*m);
    union {
        char buf[CMSG_SPACE(sizeof(int) * SA_WIRE_MAX_FDS)];
        struct cmsghdr align;
    } u;
    struct msghdr mh;
    memset(&mh, 0, sizeof(mh));
    mh.msg_iov = &iov;
    mh.msg_iovlen = 1;
    if (m->nfds && fds) {
        memset(&u, 0, sizeof(u));
        mh.msg_control = u.buf;
        mh.msg_controllen = CMSG_SPACE(sizeof(int) * m->nfds);
        struct cmsghdr* c = CMSG_FIRSTHDR(&mh);
        c->cmsg_level = SOL_SOCKET;
        c->cmsg_type = SCM_RIGHTS;
        c->cmsg_len = CMSG_LEN(sizeof(int) * m->nfds);
        memcpy(CMSG_DATA(c), fds, sizeof(int) * m->nfds);
    }
    return sendmsg(sock, &mh, MSG_NOSIGNAL) == (ssize_t)sizeof(*m) ? 0 : -1;
}

/* Returns 0, or -1 on error / end of file (the other side is gone). fds needs room for SA_WIRE_MAX_FDS. */
static inline int sa_wire_recv(int sock, sa_wire_msg* m, int* fds) {
    struct iovec iov;
    iov.iov_base = m;
    iov.iov_len = sizeof(*m);
    union {
        char buf[CMSG_SPACE(sizeof(int) * SA_WIRE_MAX_FDS)];
        struct cmsghdr align;
    } u;
    struct msghdr mh;
    memset(&mh, 0, sizeof(mh));
    memset(&u, 0, sizeof(u));
    mh.msg_iov = &iov;
    mh.msg_iovlen = 1;
    mh.msg_control = u.buf;
    mh.msg_controllen = sizeof(u.buf);
    const ssize_t n = recvmsg(sock, &mh, MSG_CMSG_CLOEXEC);
    if (n != (ssize_t)sizeof(*m)) return -1;
    uint32_t got = 0;
    for (struct cmsghdr* c = CMSG_FIRSTHDR(&mh); c; c = CMSG_NXTHDR(&mh, c)) {
        if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) {
            got = (uint32_t)((c->cmsg_len - CMSG_LEN(0)) / sizeof(int));
            if (got > SA_WIRE_MAX_FDS) got = SA_WIRE_MAX_FDS;
            memcpy(fds, CMSG_DATA(c), sizeof(int) * got);
        }
    }
    return got == m->nfds ? 0 : -1;
}
