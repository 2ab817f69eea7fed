// Host message bus: asynchronous byte messages between ranks over TCP, keyed by
// (source rank, transaction id) or delivered to a per-rank FIFO "server" channel.
//
// Same contract as the reference's native object transport inside smplib.so - smp_async_send /
// smp_async_bcast / smp_async_recv / smp_wait_recv / smp_poll_recv / smp_retrieve_object /
// smp_clean_recv_resources / smp_wait_transaction, plus the listener used by the pipeline server
// (smp_torch_has_next_message / get_next_message) - reconstructed from the Python call sites
// (reference smdistributed/modelparallel/backend/collectives.py:233-324 and
// torch/server_comm.py:162-182,286-300).  The reference rides on MPI; this one is a full TCP mesh
// bootstrapped from torchrun's rendezvous, with one receiver thread (poll) and one sender thread
// (so sends never block the training thread), and failure propagation: a rank that shuts down with
// success=0 tells every peer, which then report smpc_bus_peer_failed() (reference
// backend/core.py:165-188,227-232).  A peer whose connection ends without a goodbye frame (killed process, lost
// node) is reported the same way, so nobody waits for ever on a rank that is gone.  The listener binds the
// interface named at rendezvous, a connection has to present the job's token (exchanged through the process group)
// and a free rank, frames are size-capped and stamped with the rank of the socket they came from.
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#define SMPC_API extern "C" __attribute__((visibility("default")))

namespace {

enum Channel : uint32_t { CH_KEYED = 0, CH_SERVER = 1, CH_CONTROL = 2 };
enum Control : uint64_t { CTL_PEER_FAILED = 1, CTL_GOODBYE = 2 };

struct Frame {
  uint32_t src;
  uint32_t channel;
  uint64_t txid;
  uint64_t len;
};

struct Message {
  uint32_t src;
  uint64_t txid;
  std::vector<char> data;
};

struct Outgoing {
  int dst;
  Frame hdr;
  std::shared_ptr<std::vector<char>> payload;  // shared across bcast destinations
};

struct Bus {
  int rank = -1, size = 0;
  int listen_fd = -1;
  std::vector<int> fds;  // fds[peer], -1 for self
  std::thread rx, tx;
  std::atomic<bool> stop{false};
  std::atomic<int> failed_peer{-1};
  std::atomic<bool> io_error{false};

  std::mutex mb_mu;
  std::condition_variable mb_cv;
  std::map<std::pair<uint32_t, uint64_t>, std::deque<std::vector<char>>> mailbox;  // keyed channel
  std::deque<Message> server_q;                                                     // server channel

  std::mutex tx_mu;
  std::condition_variable tx_cv;
  std::deque<Outgoing> tx_q;
  uint64_t tx_enqueued = 0, tx_done = 0;
  std::map<uint64_t, uint64_t> tx_pending_by_id;  // txid -> outstanding sends

  std::string error;
  uint64_t token = 0;  // per-job secret every connecting peer has to present
};

constexpr uint64_t kMaxFrameBytes = 1ull << 36;  // 64 GiB: anything larger is a corrupt or hostile header

struct Hello {
  uint32_t rank;
  uint32_t magic;
  uint64_t token;
};
constexpr uint32_t kHelloMagic = 0x534d5042u;  // "SMPB"

Bus* g_bus = nullptr;
std::mutex g_mu;

bool write_all(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
    if (w < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += w;
    n -= (size_t)w;
  }
  return true;
}
bool read_all(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    ssize_t r = ::recv(fd, p, n, 0);
    if (r == 0) return false;
    if (r < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    p += r;
    n -= (size_t)r;
  }
  return true;
}

void deliver(Bus* b, const Frame& h, std::vector<char>&& data) {
  std::lock_guard<std::mutex> lk(b->mb_mu);
  if (h.channel == CH_KEYED) {
    b->mailbox[{h.src, h.txid}].emplace_back(std::move(data));
  } else if (h.channel == CH_SERVER) {
    b->server_q.push_back(Message{h.src, h.txid, std::move(data)});
  } else if (h.channel == CH_CONTROL) {
    if (h.txid == CTL_PEER_FAILED) b->failed_peer.store((int)h.src);
  }
  b->mb_cv.notify_all();
}

void mark_failed(Bus* b, int peer) {
  std::lock_guard<std::mutex> lk(b->mb_mu);
  int none = -1;
  b->failed_peer.compare_exchange_strong(none, peer);
  b->mb_cv.notify_all();
}

void rx_loop(Bus* b) {
  std::vector<pollfd> pfds;
  std::vector<int> peer_of;
  for (int p = 0; p < b->size; ++p)
    if (b->fds[p] >= 0) {
      pfds.push_back(pollfd{b->fds[p], POLLIN, 0});
      peer_of.push_back(p);
    }
  std::vector<bool> closed(pfds.size(), false), goodbye(pfds.size(), false);
  // a connection that ends, or sends something that is not a frame, before its peer said goodbye: the peer
  // is gone - wake every waiter with the failure instead of letting them block for ever
  auto drop = [&](size_t i) {
    closed[i] = true;
    pfds[i].fd = -1;  // poll ignores negative fds
    if (!goodbye[i] && !b->stop.load()) mark_failed(b, peer_of[i]);
  };
  while (!b->stop.load()) {
    int rc = ::poll(pfds.data(), pfds.size(), 50);
    if (rc <= 0) continue;
    for (size_t i = 0; i < pfds.size(); ++i) {
      if (closed[i] || !(pfds[i].revents & (POLLIN | POLLHUP | POLLERR))) continue;
      Frame h;
      if (!read_all(pfds[i].fd, &h, sizeof(h)) || h.len > kMaxFrameBytes || h.channel > CH_CONTROL) {
        drop(i);
        continue;
      }
      h.src = (uint32_t)peer_of[i];  // the socket says who is talking, not the frame
      std::vector<char> data;
      try {
        data.resize(h.len);
      } catch (const std::bad_alloc&) {
        drop(i);
        continue;
      }
      if (h.len > 0 && !read_all(pfds[i].fd, data.data(), h.len)) {
        drop(i);
        continue;
      }
      if (h.channel == CH_CONTROL && h.txid == CTL_GOODBYE) {
        goodbye[i] = true;
        continue;
      }
      deliver(b, h, std::move(data));
    }
  }
}

void tx_loop(Bus* b) {
  for (;;) {
    Outgoing o;
    {
      std::unique_lock<std::mutex> lk(b->tx_mu);
      b->tx_cv.wait(lk, [&] { return b->stop.load() || !b->tx_q.empty(); });
      if (b->tx_q.empty()) {
        if (b->stop.load()) return;
        continue;
      }
      o = std::move(b->tx_q.front());
      b->tx_q.pop_front();
    }
    bool ok = true;
    if (o.dst == b->rank) {
      std::vector<char> copy = o.payload ? *o.payload : std::vector<char>();
      deliver(b, o.hdr, std::move(copy));
    } else {
      int fd = b->fds[o.dst];
      ok = fd >= 0 && write_all(fd, &o.hdr, sizeof(o.hdr)) &&
           (o.hdr.len == 0 || write_all(fd, o.payload->data(), o.hdr.len));
    }
    if (!ok) b->io_error.store(true);
    {
      std::lock_guard<std::mutex> lk(b->tx_mu);
      b->tx_done++;
      auto it = b->tx_pending_by_id.find(o.hdr.txid);
      if (it != b->tx_pending_by_id.end() && --it->second == 0) b->tx_pending_by_id.erase(it);
    }
    b->tx_cv.notify_all();
  }
}

int enqueue(Bus* b, int dst, uint32_t channel, uint64_t txid, std::shared_ptr<std::vector<char>> payload) {
  if (dst < 0 || dst >= b->size) return -1;
  Outgoing o;
  o.dst = dst;
  o.hdr = Frame{(uint32_t)b->rank, channel, txid, payload ? (uint64_t)payload->size() : 0};
  o.payload = std::move(payload);
  {
    std::lock_guard<std::mutex> lk(b->tx_mu);
    b->tx_q.push_back(std::move(o));
    b->tx_enqueued++;
    b->tx_pending_by_id[txid]++;
  }
  b->tx_cv.notify_all();
  return 0;
}

int connect_to(const std::string& host, int port, int timeout_s) {
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  if (getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res) != 0 || !res) return -1;
  auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(timeout_s);
  int fd = -1;
  while (std::chrono::steady_clock::now() < deadline) {
    fd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (fd < 0) break;
    if (::connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
    ::close(fd);
    fd = -1;
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
  }
  freeaddrinfo(res);
  if (fd >= 0) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  }
  return fd;
}

}  // namespace

// Step 1: open a listening socket on an ephemeral port (returned), before addresses are exchanged.  ``host`` is the
// numeric IPv4 address of the interface to listen on (the one the rendezvous uses); null / empty / unparsable means
// every interface.
SMPC_API int smpc_bus_listen_on(const char* host) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_bus) return -1;
  int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) return -1;
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_ANY);
  if (host && *host) {
    in_addr a{};
    if (inet_pton(AF_INET, host, &a) == 1) addr.sin_addr = a;
  }
  addr.sin_port = 0;
  if (::bind(fd, (sockaddr*)&addr, sizeof(addr)) != 0 || ::listen(fd, 1024) != 0) {
    ::close(fd);
    return -1;  // g_bus stays null: a later attempt starts clean
  }
  socklen_t len = sizeof(addr);
  getsockname(fd, (sockaddr*)&addr, &len);
  g_bus = new Bus();
  g_bus->listen_fd = fd;
  return ntohs(addr.sin_port);
}
SMPC_API int smpc_bus_listen() { return smpc_bus_listen_on(nullptr); }

// The job's token: set after listen and before connect, the same value on every rank.
SMPC_API int smpc_bus_set_token(unsigned long long token) {
  if (!g_bus) return -1;
  g_bus->token = (uint64_t)token;
  return 0;
}

// Step 2: build the mesh.  peers = "host:port,host:port,..." indexed by rank.  Rank r connects to
// every lower rank and accepts from every higher rank.
SMPC_API int smpc_bus_connect(int rank, int size, const char* peers, int timeout_s) {
  Bus* b = g_bus;
  if (!b || b->listen_fd < 0) return -1;
  b->rank = rank;
  b->size = size;
  b->fds.assign(size, -1);
  std::vector<std::pair<std::string, int>> addr;
  std::string s(peers ? peers : "");
  size_t pos = 0;
  while (pos <= s.size() && (int)addr.size() < size) {
    size_t comma = s.find(',', pos);
    std::string item = s.substr(pos, comma == std::string::npos ? std::string::npos : comma - pos);
    size_t colon = item.rfind(':');
    if (colon == std::string::npos) return -2;
    addr.emplace_back(item.substr(0, colon), atoi(item.c_str() + colon + 1));
    if (comma == std::string::npos) break;
    pos = comma + 1;
  }
  if ((int)addr.size() != size) return -2;
  for (int p = 0; p < rank; ++p) {
    int fd = connect_to(addr[p].first, addr[p].second, timeout_s);
    if (fd < 0) return -3;
    Hello me{(uint32_t)rank, kHelloMagic, b->token};
    if (!write_all(fd, &me, sizeof(me))) return -3;
    b->fds[p] = fd;
  }
  // accept until every higher rank has presented itself; a connection with the wrong token, a rank that is out of
  // range or already connected, or one that does not complete its hello in time is closed and does not count
  auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(timeout_s);
  int missing = size - rank - 1;
  while (missing > 0) {
    auto left = std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now());
    if (left.count() <= 0) return -4;
    pollfd pf{b->listen_fd, POLLIN, 0};
    if (::poll(&pf, 1, (int)left.count()) <= 0) return -4;
    int fd = ::accept(b->listen_fd, nullptr, nullptr);
    if (fd < 0) continue;
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    timeval tv{5, 0};
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    Hello who{};
    bool ok = read_all(fd, &who, sizeof(who)) && who.magic == kHelloMagic && who.token == b->token &&
              (int)who.rank > rank && (int)who.rank < size && b->fds[who.rank] < 0;
    if (!ok) {
      ::close(fd);
      continue;
    }
    timeval none{0, 0};
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &none, sizeof(none));
    b->fds[who.rank] = fd;
    --missing;
  }
  b->rx = std::thread(rx_loop, b);
  b->tx = std::thread(tx_loop, b);
  return 0;
}

SMPC_API int smpc_bus_initialized() { return g_bus && g_bus->rank >= 0 ? 1 : 0; }
SMPC_API int smpc_bus_rank() { return g_bus ? g_bus->rank : -1; }
SMPC_API int smpc_bus_size() { return g_bus ? g_bus->size : 0; }

SMPC_API int smpc_bus_send(const void* buf, long long len, int dst, long long txid, int server) {
  if (!g_bus) return -1;
  auto payload = std::make_shared<std::vector<char>>((const char*)buf, (const char*)buf + len);
  return enqueue(g_bus, dst, server ? CH_SERVER : CH_KEYED, (uint64_t)txid, payload);
}

SMPC_API int smpc_bus_bcast(const void* buf, long long len, int n, const int* ranks, long long txid, int server) {
  if (!g_bus) return -1;
  auto payload = std::make_shared<std::vector<char>>((const char*)buf, (const char*)buf + len);
  for (int i = 0; i < n; ++i)
    if (enqueue(g_bus, ranks[i], server ? CH_SERVER : CH_KEYED, (uint64_t)txid, payload) != 0) return -1;
  return 0;
}

// Block until every send enqueued under `txid` has been written (txid < 0: all sends).
SMPC_API int smpc_bus_wait_sends(long long txid) {
  Bus* b = g_bus;
  if (!b) return -1;
  std::unique_lock<std::mutex> lk(b->tx_mu);
  if (txid < 0) {
    const uint64_t target = b->tx_enqueued;
    b->tx_cv.wait(lk, [&] { return b->tx_done >= target; });
  } else {
    b->tx_cv.wait(lk, [&] { return b->tx_pending_by_id.find((uint64_t)txid) == b->tx_pending_by_id.end(); });
  }
  return b->io_error.load() ? -2 : 0;
}

// Length of the oldest message from (src, txid), or -1 when none has arrived.
SMPC_API long long smpc_bus_poll(int src, long long txid) {
  Bus* b = g_bus;
  if (!b) return -1;
  std::lock_guard<std::mutex> lk(b->mb_mu);
  auto it = b->mailbox.find({(uint32_t)src, (uint64_t)txid});
  if (it == b->mailbox.end() || it->second.empty()) return -1;
  return (long long)it->second.front().size();
}

// Wait (timeout_ms < 0: forever) for a message from (src, txid).  Returns its length, -1 on timeout,
// -2 when a peer reported failure while waiting.
SMPC_API long long smpc_bus_wait(int src, long long txid, int timeout_ms) {
  Bus* b = g_bus;
  if (!b) return -1;
  std::unique_lock<std::mutex> lk(b->mb_mu);
  auto key = std::make_pair((uint32_t)src, (uint64_t)txid);
  auto ready = [&] {
    auto it = b->mailbox.find(key);
    return (it != b->mailbox.end() && !it->second.empty()) || b->failed_peer.load() >= 0;
  };
  if (timeout_ms < 0) {
    b->mb_cv.wait(lk, ready);
  } else if (!b->mb_cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready)) {
    return -1;
  }
  auto it = b->mailbox.find(key);
  if (it != b->mailbox.end() && !it->second.empty()) return (long long)it->second.front().size();
  return -2;
}

// Pop the oldest message from (src, txid) into out (capacity cap).  Returns bytes copied or -1.
SMPC_API long long smpc_bus_retrieve(int src, long long txid, void* out, long long cap) {
  Bus* b = g_bus;
  if (!b) return -1;
  std::lock_guard<std::mutex> lk(b->mb_mu);
  auto it = b->mailbox.find({(uint32_t)src, (uint64_t)txid});
  if (it == b->mailbox.end() || it->second.empty()) return -1;
  auto& m = it->second.front();
  if ((long long)m.size() > cap) return -1;
  memcpy(out, m.data(), m.size());
  long long n = (long long)m.size();
  it->second.pop_front();
  if (it->second.empty()) b->mailbox.erase(it);
  return n;
}

// Server channel (FIFO over all sources).  peek returns the length of the head message and fills
// src/txid, or -1; timeout_ms as above.
SMPC_API long long smpc_bus_server_peek(int* src, long long* txid, int timeout_ms) {
  Bus* b = g_bus;
  if (!b) return -1;
  std::unique_lock<std::mutex> lk(b->mb_mu);
  auto ready = [&] { return !b->server_q.empty() || b->failed_peer.load() >= 0; };
  if (timeout_ms != 0) {
    if (timeout_ms < 0)
      b->mb_cv.wait(lk, ready);
    else
      b->mb_cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready);
  }
  if (b->server_q.empty()) return b->failed_peer.load() >= 0 ? -2 : -1;
  if (src) *src = (int)b->server_q.front().src;
  if (txid) *txid = (long long)b->server_q.front().txid;
  return (long long)b->server_q.front().data.size();
}
SMPC_API long long smpc_bus_server_pop(void* out, long long cap) {
  Bus* b = g_bus;
  if (!b) return -1;
  std::lock_guard<std::mutex> lk(b->mb_mu);
  if (b->server_q.empty()) return -1;
  auto& m = b->server_q.front();
  if ((long long)m.data.size() > cap) return -1;
  memcpy(out, m.data.data(), m.data.size());
  long long n = (long long)m.data.size();
  b->server_q.pop_front();
  return n;
}

SMPC_API int smpc_bus_peer_failed() { return g_bus ? g_bus->failed_peer.load() : -1; }

// Coordinated shutdown: success=0 notifies every peer first so they can abort instead of hanging.
SMPC_API int smpc_bus_shutdown(int success) {
  std::lock_guard<std::mutex> lk(g_mu);
  Bus* b = g_bus;
  if (!b) return 0;
  if (b->rank >= 0) {
    for (int p = 0; p < b->size; ++p)
      if (p != b->rank) enqueue(b, p, CH_CONTROL, success ? CTL_GOODBYE : CTL_PEER_FAILED, nullptr);
    {
      std::unique_lock<std::mutex> tl(b->tx_mu);
      const uint64_t target = b->tx_enqueued;
      b->tx_cv.wait_for(tl, std::chrono::seconds(5), [&] { return b->tx_done >= target; });
    }
    {
      // under the sender's mutex: a sender that has just found its queue empty must either see the flag or be
      // asleep when the notification comes - set outside the lock, the wake-up could fall between the two and
      // the join below would wait for ever
      std::lock_guard<std::mutex> tl(b->tx_mu);
      b->stop.store(true);
    }
    b->tx_cv.notify_all();
    if (b->tx.joinable()) b->tx.join();
    // everything this rank had to say is on the wire: end the connections, which also wakes the receiver should it
    // be in the middle of a frame whose sender has stalled, then wait for it
    for (int fd : b->fds)
      if (fd >= 0) ::shutdown(fd, SHUT_RDWR);
    if (b->rx.joinable()) b->rx.join();
    for (int fd : b->fds)
      if (fd >= 0) ::close(fd);
  }
  if (b->listen_fd >= 0) ::close(b->listen_fd);
  delete b;
  g_bus = nullptr;
  return 0;
}
