// MPI shim runtime: full mesh of AF_UNIX stream socketpairs inherited from bin/jb_mpirun.
// Environment set by the launcher: JB_MPI_RANK, JB_MPI_SIZE, JB_MPI_FDS="fd,fd,...".
// Point-to-point = framed messages {tag, bytes} on the pair's socket; a sender that cannot write
// keeps draining every peer into per-peer inboxes, so simultaneous large exchanges (the reference's
// MPI_Sendrecv_replace ring of n^2/p doubles, main.cpp:639) cannot deadlock.
#include "mpi.h"

#include <errno.h>
#include <fcntl.h>
#include <poll.h>
#include <sys/socket.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <deque>
#include <string>
#include <vector>

namespace {

struct Header {
  int32_t tag;
  int32_t pad;
  int64_t bytes;
};

struct Peer {
  int fd = -1;
  std::string inbox;  // raw bytes received but not yet consumed
  size_t inpos = 0;
  std::deque<int> fds;  // descriptors that arrived as SCM_RIGHTS ancillary data (jb_mpi_recv_fd)
};

int g_rank = 0, g_size = 1;
bool g_init = false;
std::vector<Peer> g_peers;

struct DType {
  size_t extent;
  bool builtin;
};
std::vector<DType> g_types;
std::vector<MPI_User_function*> g_ops;
const int kUserOpBase = 100;

[[noreturn]] void die(const char* what) {
  fprintf(stderr, "[jb-mpi rank %d] fatal: %s (errno %d: %s)\n", g_rank, what, errno, strerror(errno));
  _exit(70);
}

size_t type_size(MPI_Datatype t) {
  switch (t) {
    case MPI_BYTE: case MPI_CHAR: return 1;
    case MPI_INT: case MPI_UNSIGNED: case MPI_FLOAT: return 4;
    case MPI_LONG_LONG: case MPI_DOUBLE: return 8;
    default:
      if (t >= 1000 && (size_t)(t - 1000) < g_types.size()) return g_types[t - 1000].extent;
      die("unknown datatype");
  }
}

void compact(Peer& p) {
  if (p.inpos > (1u << 20) && p.inpos * 2 > p.inbox.size()) {
    p.inbox.erase(0, p.inpos);
    p.inpos = 0;
  }
}

// recv() that also collects file descriptors passed along with the bytes (SCM_RIGHTS).  Every read
// of a peer socket goes through here: a plain recv() would silently drop (and close) descriptors
// attached to the bytes it returns.
ssize_t recv_with_fds(Peer& p, void* buf, size_t len, int flags) {
  iovec iov;
  iov.iov_base = buf;
  iov.iov_len = len;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(8 * sizeof(int))];
  msghdr mh;
  memset(&mh, 0, sizeof(mh));
  mh.msg_iov = &iov;
  mh.msg_iovlen = 1;
  mh.msg_control = ctrl;
  mh.msg_controllen = sizeof(ctrl);
  ssize_t r = recvmsg(p.fd, &mh, flags | MSG_CMSG_CLOEXEC);
  if (r > 0)
    for (cmsghdr* c = CMSG_FIRSTHDR(&mh); c; c = CMSG_NXTHDR(&mh, c))
      if (c->cmsg_level == SOL_SOCKET && c->cmsg_type == SCM_RIGHTS) {
        const size_t nfd = (c->cmsg_len - CMSG_LEN(0)) / sizeof(int);
        int got[8];
        memcpy(got, CMSG_DATA(c), nfd * sizeof(int));
        for (size_t i = 0; i < nfd; ++i) p.fds.push_back(got[i]);
      }
  return r;
}

// read whatever is available on fd into the peer's inbox (non-blocking)
bool drain(Peer& p) {
  char buf[1 << 16];
  bool any = false;
  for (;;) {
    ssize_t r = recv_with_fds(p, buf, sizeof(buf), MSG_DONTWAIT);
    if (r > 0) {
      p.inbox.append(buf, (size_t)r);
      any = true;
      if ((size_t)r < sizeof(buf)) break;
    } else if (r == 0) {
      break;  // peer closed
    } else {
      if (errno == EAGAIN || errno == EWOULDBLOCK) break;
      if (errno == EINTR) continue;
      die("recv");
    }
  }
  return any;
}

void send_all(int dest, const void* data, size_t len) {
  const char* p = (const char*)data;
  Peer& d = g_peers[dest];
  while (len) {
    ssize_t w = send(d.fd, p, len, MSG_DONTWAIT | MSG_NOSIGNAL);
    if (w > 0) {
      p += w;
      len -= (size_t)w;
      continue;
    }
    if (w < 0 && errno == EINTR) continue;
    if (w < 0 && errno != EAGAIN && errno != EWOULDBLOCK) die("send");
    // cannot write now: wait for writability, draining everybody meanwhile
    std::vector<pollfd> pf;
    for (int r = 0; r < g_size; ++r) {
      if (r == g_rank) continue;
      pollfd e;
      e.fd = g_peers[r].fd;
      e.events = POLLIN | (r == dest ? POLLOUT : 0);
      e.revents = 0;
      pf.push_back(e);
    }
    if (poll(pf.data(), pf.size(), -1) < 0 && errno != EINTR) die("poll");
    for (int r = 0; r < g_size; ++r)
      if (r != g_rank) drain(g_peers[r]);
  }
}

void recv_exact(int src, void* out, size_t len) {
  Peer& s = g_peers[src];
  char* o = (char*)out;
  while (len) {
    size_t have = s.inbox.size() - s.inpos;
    if (have) {
      size_t take = have < len ? have : len;
      memcpy(o, s.inbox.data() + s.inpos, take);
      s.inpos += take;
      o += take;
      len -= take;
      if (s.inpos == s.inbox.size()) {
        s.inbox.clear();
        s.inpos = 0;
      } else {
        compact(s);
      }
      continue;
    }
    // large payloads: read straight into the user buffer
    ssize_t r = recv_with_fds(s, o, len, 0);
    if (r > 0) {
      o += r;
      len -= (size_t)r;
    } else if (r == 0) {
      die("peer closed the connection (a rank died?)");
    } else if (errno != EINTR) {
      die("recv");
    }
  }
}

void p2p_send(const void* buf, size_t bytes, int dest, int tag) {
  if (dest == g_rank) die("send to self is not supported");
  Header h;
  h.tag = tag;
  h.pad = 0;
  h.bytes = (int64_t)bytes;
  send_all(dest, &h, sizeof(h));
  send_all(dest, buf, bytes);
}

void p2p_recv(void* buf, size_t bytes, int src, int tag, MPI_Status* st) {
  Header h;
  recv_exact(src, &h, sizeof(h));
  if ((size_t)h.bytes > bytes) die("message truncated");
  if (tag != MPI_ANY_TAG && h.tag != tag) die("tag mismatch (the shim delivers in order per pair)");
  recv_exact(src, buf, (size_t)h.bytes);
  if (st) {
    st->MPI_SOURCE = src;
    st->MPI_TAG = h.tag;
    st->MPI_ERROR = 0;
    st->count_bytes = (int)h.bytes;
  }
}

template <typename T>
void builtin_reduce(T* inout, const T* in, int n, MPI_Op op) {
  for (int i = 0; i < n; ++i) {
    if (op == MPI_SUM) inout[i] = in[i] + inout[i];
    else if (op == MPI_MAX) inout[i] = in[i] > inout[i] ? in[i] : inout[i];
    else if (op == MPI_MIN) inout[i] = in[i] < inout[i] ? in[i] : inout[i];
    else die("unknown reduction op");
  }
}

// inout = in o inout   (MPI's convention for user functions)
void apply_op(void* in, void* inout, int count, MPI_Datatype type, MPI_Op op) {
  if (op >= kUserOpBase) {
    MPI_User_function* fn = g_ops[op - kUserOpBase];
    if (!fn) die("freed op used");
    fn(in, inout, &count, &type);
    return;
  }
  switch (type) {
    case MPI_INT: builtin_reduce((int*)inout, (const int*)in, count, op); break;
    case MPI_UNSIGNED: builtin_reduce((unsigned*)inout, (const unsigned*)in, count, op); break;
    case MPI_LONG_LONG: builtin_reduce((long long*)inout, (const long long*)in, count, op); break;
    case MPI_DOUBLE: builtin_reduce((double*)inout, (const double*)in, count, op); break;
    case MPI_FLOAT: builtin_reduce((float*)inout, (const float*)in, count, op); break;
    default: die("built-in reduction on unsupported datatype");
  }
}

}  // namespace

extern "C" {

int MPI_Init(int*, char***) {
  const char* r = getenv("JB_MPI_RANK");
  const char* s = getenv("JB_MPI_SIZE");
  const char* f = getenv("JB_MPI_FDS");
  if (r && s && f) {
    g_rank = atoi(r);
    g_size = atoi(s);
    g_peers.assign(g_size, Peer());
    const char* p = f;
    for (int i = 0; i < g_size; ++i) {
      g_peers[i].fd = atoi(p);
      const char* c = strchr(p, ',');
      p = c ? c + 1 : p + strlen(p);
    }
  } else {
    g_rank = 0;
    g_size = 1;
    g_peers.assign(1, Peer());
  }
  setvbuf(stdout, nullptr, _IOLBF, 0);  // report lines appear as they are printed (SURVEY App. D v)
  g_init = true;
  return MPI_SUCCESS;
}

int MPI_Finalize(void) {
  fflush(stdout);
  if (g_size > 1) MPI_Barrier(MPI_COMM_WORLD);
  for (auto& p : g_peers)
    if (p.fd >= 0) close(p.fd);
  g_init = false;
  return MPI_SUCCESS;
}

int MPI_Abort(MPI_Comm, int code) {
  fflush(stdout);
  _exit(code ? code : 1);
}

int MPI_Comm_size(MPI_Comm, int* size) { *size = g_size; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm, int* rank) { *rank = g_rank; return MPI_SUCCESS; }

double MPI_Wtime(void) {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm) {
  p2p_send(buf, (size_t)count * type_size(type), dest, tag);
  return MPI_SUCCESS;
}

int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm, MPI_Status* st) {
  p2p_recv(buf, (size_t)count * type_size(type), source, tag, st);
  return MPI_SUCCESS;
}

int MPI_Sendrecv_replace(void* buf, int count, MPI_Datatype type, int dest, int sendtag, int source,
                         int recvtag, MPI_Comm, MPI_Status* st) {
  const size_t bytes = (size_t)count * type_size(type);
  if (g_size == 1 || (dest == g_rank && source == g_rank)) return MPI_SUCCESS;
  // send_all drains incoming traffic while it waits, so everybody may send first (dest and source
  // can even be the same endpoint, as happens for p = 2)
  std::vector<char> tmp((const char*)buf, (const char*)buf + bytes);
  p2p_send(tmp.data(), bytes, dest, sendtag);
  p2p_recv(buf, bytes, source, recvtag, st);
  return MPI_SUCCESS;
}

int MPI_Bcast(void* buf, int count, MPI_Datatype type, int root, MPI_Comm) {
  const size_t bytes = (size_t)count * type_size(type);
  if (g_size == 1) return MPI_SUCCESS;
  if (g_rank == root) {
    for (int r = 0; r < g_size; ++r)
      if (r != root) p2p_send(buf, bytes, r, -7);
  } else {
    p2p_recv(buf, bytes, root, -7, nullptr);
  }
  return MPI_SUCCESS;
}

int MPI_Barrier(MPI_Comm) {
  int x = 0, y = 0;
  return MPI_Allreduce(&x, &y, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD);
}

int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype type, MPI_Op op,
                  MPI_Comm) {
  const size_t bytes = (size_t)count * type_size(type);
  if (recvbuf != sendbuf) memcpy(recvbuf, sendbuf, bytes);
  if (g_size == 1) return MPI_SUCCESS;
  if (g_rank == 0) {
    // fold strictly in rank order: acc = r0; acc = acc o r_k  (inout = in o inout with in = acc).
    // This is what makes ties of the reference's non-commutative MINPIV go to the highest rank.
    std::vector<char> acc((const char*)recvbuf, (const char*)recvbuf + bytes), cur(bytes);
    for (int r = 1; r < g_size; ++r) {
      p2p_recv(cur.data(), bytes, r, -8, nullptr);
      apply_op(acc.data(), cur.data(), count, type, op);
      acc.swap(cur);
    }
    memcpy(recvbuf, acc.data(), bytes);
  } else {
    p2p_send(recvbuf, bytes, 0, -8);
  }
  return MPI_Bcast(recvbuf, count, type, 0, MPI_COMM_WORLD);
}

int MPI_Gather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf, int, MPI_Datatype,
               int root, MPI_Comm) {
  const size_t bytes = (size_t)sendcount * type_size(sendtype);
  if (g_rank == root) {
    for (int r = 0; r < g_size; ++r) {
      char* dst = (char*)recvbuf + (size_t)r * bytes;
      if (r == root) memcpy(dst, sendbuf, bytes);
      else p2p_recv(dst, bytes, r, -9, nullptr);
    }
  } else {
    p2p_send(sendbuf, bytes, root, -9);
  }
  return MPI_SUCCESS;
}

int MPI_Allgather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf,
                  int recvcount, MPI_Datatype recvtype, MPI_Comm) {
  MPI_Gather(sendbuf, sendcount, sendtype, recvbuf, recvcount, recvtype, 0, MPI_COMM_WORLD);
  return MPI_Bcast(recvbuf, recvcount * g_size, recvtype, 0, MPI_COMM_WORLD);
}

// ---- extension (not MPI): pass an open file descriptor to another rank of the same host.  Used to
// share CUDA VMM allocations between the rank processes (csrc/comm/symm_heap.cpp).  The descriptor
// rides as SCM_RIGHTS ancillary data on the one payload byte of an ordinary framed message, so it
// stays in order with every other message of the pair.
int jb_mpi_send_fd(int fd, int dest) {
  if (dest == g_rank || dest < 0 || dest >= g_size) return -1;
  Header h;
  h.tag = -12;
  h.pad = 0;
  h.bytes = 1;
  send_all(dest, &h, sizeof(h));
  char byte = 'F';
  iovec iov;
  iov.iov_base = &byte;
  iov.iov_len = 1;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
  memset(ctrl, 0, sizeof(ctrl));
  msghdr mh;
  memset(&mh, 0, sizeof(mh));
  mh.msg_iov = &iov;
  mh.msg_iovlen = 1;
  mh.msg_control = ctrl;
  mh.msg_controllen = sizeof(ctrl);
  cmsghdr* c = CMSG_FIRSTHDR(&mh);
  c->cmsg_level = SOL_SOCKET;
  c->cmsg_type = SCM_RIGHTS;
  c->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(c), &fd, sizeof(int));
  for (;;) {
    ssize_t w = sendmsg(g_peers[dest].fd, &mh, MSG_NOSIGNAL);  // blocking: one byte
    if (w == 1) return 0;
    if (w < 0 && (errno == EINTR)) continue;
    if (w < 0 && (errno == EAGAIN || errno == EWOULDBLOCK)) {
      for (int r = 0; r < g_size; ++r)
        if (r != g_rank) drain(g_peers[r]);
      continue;
    }
    die("sendmsg(SCM_RIGHTS)");
  }
}

int jb_mpi_recv_fd(int src) {
  if (src == g_rank || src < 0 || src >= g_size) return -1;
  char byte = 0;
  p2p_recv(&byte, 1, src, -12, nullptr);
  Peer& p = g_peers[src];
  if (p.fds.empty()) return -1;
  int fd = p.fds.front();
  p.fds.pop_front();
  return fd;
}

int MPI_Address(void* location, MPI_Aint* address) { *address = (MPI_Aint)location; return MPI_SUCCESS; }
int MPI_Get_address(const void* location, MPI_Aint* address) { *address = (MPI_Aint)location; return MPI_SUCCESS; }

int MPI_Type_struct(int count, int* blocklens, MPI_Aint* displs, MPI_Datatype* types, MPI_Datatype* newtype) {
  size_t end = 0, align = 1;
  for (int i = 0; i < count; ++i) {
    size_t sz = type_size(types[i]);
    size_t e = (size_t)displs[i] + sz * (size_t)blocklens[i];
    if (e > end) end = e;
    if (sz > align && sz <= 8) align = sz;
  }
  DType t;
  t.extent = (end + align - 1) / align * align;  // {double,int,int,int} -> 24 bytes
  t.builtin = false;
  g_types.push_back(t);
  *newtype = 1000 + (int)g_types.size() - 1;
  return MPI_SUCCESS;
}
int MPI_Type_commit(MPI_Datatype*) { return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype* t) { *t = 0; return MPI_SUCCESS; }

int MPI_Op_create(MPI_User_function* fn, int, MPI_Op* op) {
  g_ops.push_back(fn);
  *op = kUserOpBase + (int)g_ops.size() - 1;
  return MPI_SUCCESS;
}
int MPI_Op_free(MPI_Op* op) {
  if (*op >= kUserOpBase) g_ops[*op - kUserOpBase] = nullptr;
  *op = 0;
  return MPI_SUCCESS;
}

}  // extern "C"
