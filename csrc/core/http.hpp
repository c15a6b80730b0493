// http.hpp — minimal HTTP/1.1 server (thread per connection, route patterns with :params, JSON helpers) and a
// blocking client. Serves the master's REST API; long-poll endpoints replace the reference's WebSockets
// (master/internal/api/*.go, core.go:313-606 register the reference's echo routes).
#pragma once
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cstring>
#include <functional>
#include <map>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include <netdb.h>
#include <poll.h>

#include "json.hpp"
#include "tls.hpp"

namespace det {

struct HttpRequest {
  std::string method, path, query_string, body;
  std::string remote_addr;   // peer IPv4 address of the connection
  std::map<std::string, std::string> headers, query, params;
  Json json() const { return body.empty() ? Json::object() : Json::parse(body); }
  std::string q(const std::string& k, const std::string& def = "") const {
    auto it = query.find(k);
    return it == query.end() ? def : it->second;
  }
  long long qi(const std::string& k, long long def = 0) const {
    auto it = query.find(k);
    return it == query.end() || it->second.empty() ? def : atoll(it->second.c_str());
  }
};

struct HttpResponse {
  int status = 200;
  std::string content_type = "application/json";
  std::string body;
  // When set, the server hands the client socket to this function instead of writing a response (protocol upgrades:
  // the handler speaks for itself and returns when the conversation is over; the server closes the socket afterwards).
  std::function<void(int)> hijack;
  static HttpResponse json(const Json& j, int status = 200) { HttpResponse r; r.status = status; r.body = j.dump(); return r; }
  static HttpResponse error(int status, const std::string& msg) {
    Json j = Json::object();
    j["error"] = msg;
    j["code"] = status;
    return json(j, status);
  }
  static HttpResponse text(const std::string& s, int status = 200, const std::string& ct = "text/plain") {
    HttpResponse r; r.status = status; r.content_type = ct; r.body = s; return r;
  }
};

inline std::string url_decode(const std::string& s) {
  std::string o;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '%' && i + 2 < s.size()) { o.push_back(static_cast<char>(strtol(s.substr(i + 1, 2).c_str(), nullptr, 16))); i += 2; }
    else if (s[i] == '+') o.push_back(' ');
    else o.push_back(s[i]);
  }
  return o;
}

class HttpServer {
 public:
  using Handler = std::function<HttpResponse(const HttpRequest&)>;
  void route(const std::string& method, const std::string& pattern, Handler h) { routes_.push_back(Route{method, split(pattern), std::move(h)}); }
  void set_middleware(std::function<bool(const HttpRequest&, HttpResponse&)> mw) { middleware_ = std::move(mw); }
  // Serve HTTPS instead of HTTP (PEM certificate chain + key). Call before listen(); returns false and fills *err when the
  // files cannot be loaded.
  bool enable_tls(const std::string& cert_file, const std::string& key_file, std::string* err) {
    tls_ctx_ = tls_server_context(cert_file, key_file, err);
    return tls_ctx_ != nullptr;
  }
  bool tls_enabled() const { return tls_ctx_ != nullptr; }
  void set_cors(bool on) { cors_ = on; }

  // Binds and listens; returns the bound port (useful with port 0) or -1.
  int listen(const std::string& host, int port) {
    fd_ = ::socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);   // never inherited by the harness processes the master forks
    if (fd_ < 0) return -1;
    int one = 1;
    setsockopt(fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons(static_cast<uint16_t>(port));
    addr.sin_addr.s_addr = host.empty() || host == "0.0.0.0" ? INADDR_ANY : inet_addr(host.c_str());
    if (::bind(fd_, reinterpret_cast<sockaddr*>(&addr), sizeof addr) < 0) { ::close(fd_); fd_ = -1; return -1; }
    if (::listen(fd_, 256) < 0) { ::close(fd_); fd_ = -1; return -1; }
    socklen_t len = sizeof addr;
    getsockname(fd_, reinterpret_cast<sockaddr*>(&addr), &len);
    port_ = ntohs(addr.sin_port);
    return port_;
  }
  void serve_forever() {
    while (!stop_) {
      sockaddr_in peer{};
      socklen_t plen = sizeof peer;
      int c = ::accept4(fd_, reinterpret_cast<sockaddr*>(&peer), &plen, SOCK_CLOEXEC);
      if (c < 0) { if (stop_) break; continue; }
      int one = 1;
      setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
      char ip[INET_ADDRSTRLEN] = {0};
      inet_ntop(AF_INET, &peer.sin_addr, ip, sizeof ip);
      std::thread([this, c, addr = std::string(ip)] { handle(c, addr); }).detach();
    }
  }
  void start() { thread_ = std::thread([this] { serve_forever(); }); }
  void stop() {
    stop_ = true;
    if (fd_ >= 0) { ::shutdown(fd_, SHUT_RDWR); ::close(fd_); fd_ = -1; }
    if (thread_.joinable()) thread_.join();
  }
  int port() const { return port_; }

 private:
  struct Route { std::string method; std::vector<std::string> parts; Handler h; };
  static std::vector<std::string> split(const std::string& p) {
    std::vector<std::string> out;
    std::stringstream ss(p);
    std::string item;
    while (std::getline(ss, item, '/')) if (!item.empty()) out.push_back(item);
    return out;
  }
  static bool read_request(Stream& c, HttpRequest& req) {
    std::string buf;
    char tmp[8192];
    size_t header_end = std::string::npos;
    while (header_end == std::string::npos) {
      ssize_t n = c.read(tmp, sizeof tmp);
      if (n <= 0) return false;
      buf.append(tmp, static_cast<size_t>(n));
      header_end = buf.find("\r\n\r\n");
      if (buf.size() > (1u << 20) && header_end == std::string::npos) return false;
    }
    std::istringstream hs(buf.substr(0, header_end));
    std::string line;
    std::getline(hs, line);
    if (!line.empty() && line.back() == '\r') line.pop_back();
    std::istringstream rl(line);
    std::string target, ver;
    rl >> req.method >> target >> ver;
    auto qpos = target.find('?');
    req.path = url_decode(target.substr(0, qpos));
    if (qpos != std::string::npos) {
      req.query_string = target.substr(qpos + 1);
      std::stringstream qs(req.query_string);
      std::string kv;
      while (std::getline(qs, kv, '&')) {
        auto eq = kv.find('=');
        req.query[url_decode(kv.substr(0, eq))] = eq == std::string::npos ? "" : url_decode(kv.substr(eq + 1));
      }
    }
    while (std::getline(hs, line)) {
      if (!line.empty() && line.back() == '\r') line.pop_back();
      auto colon = line.find(':');
      if (colon == std::string::npos) continue;
      std::string k = line.substr(0, colon), v = line.substr(colon + 1);
      for (auto& ch : k) ch = static_cast<char>(tolower(ch));
      while (!v.empty() && v.front() == ' ') v.erase(v.begin());
      req.headers[k] = v;
    }
    size_t content_len = 0;
    auto it = req.headers.find("content-length");
    if (it != req.headers.end()) {
      // digits only, and bounded: a model-definition upload is capped at 96 MB by the clients; nothing legitimate is larger
      const std::string& v = it->second;
      if (v.empty() || v.size() > 10 || v.find_first_not_of("0123456789") != std::string::npos) return false;
      const long long n = atoll(v.c_str());
      if (n < 0 || n > (256LL << 20)) return false;
      content_len = static_cast<size_t>(n);
    }
    req.body = buf.substr(header_end + 4);
    while (req.body.size() < content_len) {
      ssize_t n = c.read(tmp, sizeof tmp);
      if (n <= 0) return false;
      req.body.append(tmp, static_cast<size_t>(n));
    }
    return true;
  }
  void handle(int cfd, const std::string& remote_addr) {
    Stream c(cfd);
    if (tls_ctx_ && !c.accept_tls(tls_ctx_)) return;   // not a TLS client (or a failed handshake): drop the connection
    HttpRequest req;
    req.remote_addr = remote_addr;
    if (read_request(c, req)) {
      HttpResponse resp = dispatch(req);
      if (resp.hijack) {
        if (!c.tls()) {
          resp.hijack(c.fd());
        } else {
          // the handler speaks plain bytes on a socket of its own; pump them through the TLS session
          int sv[2];
          if (::socketpair(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0, sv) == 0) {
            std::thread pump([&c, fd = sv[0]] {
              pollfd fds[2] = {{c.fd(), POLLIN, 0}, {fd, POLLIN, 0}};
              char buf[16384];
              for (;;) {
                if (!c.pending() && ::poll(fds, 2, 3600 * 1000) <= 0) break;
                if (c.pending() || (fds[0].revents & POLLIN)) {
                  const ssize_t n = c.read(buf, sizeof buf);
                  if (n <= 0 || ::send(fd, buf, static_cast<size_t>(n), MSG_NOSIGNAL) <= 0) break;
                }
                if (fds[1].revents & POLLIN) {
                  const ssize_t n = ::recv(fd, buf, sizeof buf, 0);
                  if (n <= 0 || !c.write_all(buf, static_cast<size_t>(n))) break;
                }
                if ((fds[0].revents | fds[1].revents) & (POLLHUP | POLLERR | POLLNVAL)) break;
              }
              ::shutdown(fd, SHUT_RDWR);
            });
            resp.hijack(sv[1]);
            ::shutdown(sv[1], SHUT_RDWR);
            pump.join();
            ::close(sv[0]);
            ::close(sv[1]);
          }
        }
        return;
      }
      static const std::map<int, const char*> reason = {{200, "OK"}, {201, "Created"}, {204, "No Content"}, {400, "Bad Request"},
                                                        {401, "Unauthorized"}, {403, "Forbidden"}, {404, "Not Found"},
                                                        {409, "Conflict"}, {500, "Internal Server Error"}};
      auto rit = reason.find(resp.status);
      std::ostringstream os;
      os << "HTTP/1.1 " << resp.status << " " << (rit == reason.end() ? "Status" : rit->second) << "\r\n"
         << "Content-Type: " << resp.content_type << "\r\nContent-Length: " << resp.body.size();
      if (cors_) {   // enable_cors: other origins (a WebUI served elsewhere) may call the API with a bearer token
        auto origin = req.headers.find("origin");
        os << "\r\nAccess-Control-Allow-Origin: " << (origin == req.headers.end() ? std::string("*") : origin->second)
           << "\r\nAccess-Control-Allow-Methods: GET, POST, PUT, PATCH, DELETE, OPTIONS"
           << "\r\nAccess-Control-Allow-Headers: Authorization, Content-Type, X-Requested-With"
           << "\r\nAccess-Control-Allow-Credentials: true\r\nVary: Origin";
      }
      os << "\r\nConnection: close\r\n\r\n" << resp.body;
      const std::string out = os.str();
      c.write_all(out.data(), out.size());
    }
  }
  HttpResponse dispatch(HttpRequest& req) {
    try {
      if (req.method == "OPTIONS") return HttpResponse::text("", 204);
      auto parts = split(req.path);
      bool path_matched = false;
      for (auto& r : routes_) {
        if (r.parts.size() != parts.size() && !(r.parts.size() <= parts.size() && !r.parts.empty() && r.parts.back() == "*")) continue;
        std::map<std::string, std::string> params;
        bool ok = true;
        for (size_t i = 0; i < r.parts.size() && ok; ++i) {
          if (r.parts[i] == "*") { std::string rest; for (size_t j = i; j < parts.size(); ++j) rest += (j > i ? "/" : "") + parts[j]; params["*"] = rest; break; }
          if (r.parts[i][0] == ':') params[r.parts[i].substr(1)] = parts[i];
          else if (r.parts[i] != parts[i]) ok = false;
        }
        if (!ok) continue;
        path_matched = true;
        if (r.method != req.method) continue;
        req.params = params;
        HttpResponse resp;
        if (middleware_ && !middleware_(req, resp)) return resp;
        return r.h(req);
      }
      return HttpResponse::error(path_matched ? 405 : 404, path_matched ? "method not allowed" : "no route for " + req.path);
    } catch (const std::exception& e) {
      return HttpResponse::error(400, e.what());
    }
  }
  std::vector<Route> routes_;
  std::function<bool(const HttpRequest&, HttpResponse&)> middleware_;
  int fd_ = -1, port_ = 0;
  SSL_CTX* tls_ctx_ = nullptr;
  bool cors_ = false;
  std::atomic<bool> stop_{false};
  std::thread thread_;
};

// Dotted quad, resolvable name, or — for this machine's own host name when the resolver does not know it — loopback.
inline in_addr_t resolve_ipv4(const std::string& host) {
  if (host == "localhost") return inet_addr("127.0.0.1");
  in_addr_t a = inet_addr(host.c_str());
  if (a != INADDR_NONE) return a;
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  if (getaddrinfo(host.c_str(), nullptr, &hints, &res) == 0 && res != nullptr) {
    a = reinterpret_cast<sockaddr_in*>(res->ai_addr)->sin_addr.s_addr;
    freeaddrinfo(res);
    return a;
  }
  char self[256] = {0};
  if (gethostname(self, sizeof self - 1) == 0 && host == self) return inet_addr("127.0.0.1");
  return INADDR_NONE;
}

// Blocking one-shot client: returns status (or -1) and fills body. With tls.enabled the request goes over TLS and the
// server's certificate is checked against tls.ca_file / tls.cert_name (see tls.hpp).
inline int http_request(const std::string& host, int port, const std::string& method, const std::string& path,
                        const std::string& body, std::string* out, int timeout_s = 60, const std::string& auth = "",
                        const TlsClientOptions& tls = TlsClientOptions()) {
  int fd = ::socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (fd < 0) return -1;
  timeval tv{timeout_s, 0};
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  addr.sin_addr.s_addr = resolve_ipv4(host);
  if (::connect(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) < 0) { ::close(fd); return -1; }
  Stream c(fd);
  if (tls.enabled) {
    std::string err;
    SSL_CTX* ctx = tls_client_context(tls, &err);
    if (ctx == nullptr || !c.connect_tls(ctx, host, tls, &err)) {
      if (out) *out = err;
      return -1;
    }
  }
  std::ostringstream os;
  os << method << " " << path << " HTTP/1.1\r\nHost: " << host << "\r\nContent-Type: application/json\r\nContent-Length: " << body.size()
     << "\r\nConnection: close\r\n";
  if (!auth.empty()) os << "Authorization: Bearer " << auth << "\r\n";
  os << "\r\n" << body;
  const std::string req = os.str();
  if (!c.write_all(req.data(), req.size())) return -1;
  std::string buf;
  char tmp[8192];
  ssize_t n;
  while ((n = c.read(tmp, sizeof tmp)) > 0) buf.append(tmp, static_cast<size_t>(n));
  auto he = buf.find("\r\n\r\n");
  if (he == std::string::npos) return -1;
  int status = -1;
  sscanf(buf.c_str(), "HTTP/%*s %d", &status);
  if (out) *out = buf.substr(he + 4);
  return status;
}

// Reverse-proxy helper: replay one request against host:port and hand back status / content type / body.
inline HttpResponse http_forward(const std::string& host, int port, const HttpRequest& req, const std::string& path, int timeout_s = 60) {
  HttpResponse out;
  int fd = ::socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (fd < 0) return HttpResponse::error(502, "proxy: socket() failed");
  timeval tv{timeout_s, 0};
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  addr.sin_addr.s_addr = resolve_ipv4(host);
  if (::connect(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) < 0) { ::close(fd); return HttpResponse::error(502, "proxy: task is not accepting connections"); }
  std::ostringstream os;
  os << req.method << " " << path << " HTTP/1.1\r\nHost: " << host << ":" << port << "\r\nConnection: close\r\nContent-Length: " << req.body.size() << "\r\n";
  for (auto& kv : req.headers) {
    std::string k = kv.first;
    for (auto& ch : k) ch = static_cast<char>(tolower(ch));
    if (k == "host" || k == "connection" || k == "content-length" || k == "authorization") continue;
    os << kv.first << ": " << kv.second << "\r\n";
  }
  os << "\r\n" << req.body;
  const std::string wire = os.str();
  if (::send(fd, wire.data(), wire.size(), MSG_NOSIGNAL) < 0) { ::close(fd); return HttpResponse::error(502, "proxy: send failed"); }
  std::string buf;
  char tmp[16384];
  ssize_t n;
  while ((n = ::recv(fd, tmp, sizeof tmp, 0)) > 0) buf.append(tmp, static_cast<size_t>(n));
  ::close(fd);
  const auto he = buf.find("\r\n\r\n");
  if (he == std::string::npos) return HttpResponse::error(502, "proxy: malformed response from task");
  int status = 502;
  sscanf(buf.c_str(), "HTTP/%*s %d", &status);
  out.status = status;
  out.body = buf.substr(he + 4);
  out.content_type = "application/octet-stream";
  std::string head = buf.substr(0, he);
  std::string lower = head;
  for (auto& ch : lower) ch = static_cast<char>(tolower(ch));
  const auto ct = lower.find("\r\ncontent-type:");
  if (ct != std::string::npos) {
    const size_t b = ct + 15, e = head.find("\r\n", b);
    std::string v = head.substr(b, e == std::string::npos ? std::string::npos : e - b);
    while (!v.empty() && v.front() == ' ') v.erase(v.begin());
    out.content_type = v;
  }
  if (lower.find("transfer-encoding: chunked") != std::string::npos) {   // de-chunk: our own reply carries Content-Length
    std::string body;
    size_t pos = 0;
    while (pos < out.body.size()) {
      const size_t nl = out.body.find("\r\n", pos);
      if (nl == std::string::npos) break;
      const size_t len = strtoul(out.body.substr(pos, nl - pos).c_str(), nullptr, 16);
      if (len == 0) break;
      body.append(out.body, nl + 2, len);
      pos = nl + 2 + len + 2;
    }
    out.body = body;
  }
  return out;
}

// Protocol-upgrade tunnel (WebSocket): replay the upgrade request against host:port, then copy bytes both ways between
// the client socket and the task until either side closes or nothing moves for `idle_s` seconds. The 101 response is the
// task's own; the master does not interpret frames.
inline void http_tunnel(const std::string& host, int port, const HttpRequest& req, const std::string& path, int client, int idle_s = 3600) {
  auto send_all = [](int fd, const char* p, size_t n) {
    while (n > 0) {
      ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
      if (k <= 0) return false;
      p += k;
      n -= static_cast<size_t>(k);
    }
    return true;
  };
  int fd = ::socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  addr.sin_addr.s_addr = resolve_ipv4(host);
  if (fd < 0 || ::connect(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) < 0) {
    static const char bad[] = "HTTP/1.1 502 Bad Gateway\r\nContent-Length: 0\r\nConnection: close\r\n\r\n";
    send_all(client, bad, sizeof bad - 1);
    if (fd >= 0) ::close(fd);
    return;
  }
  std::ostringstream os;
  os << req.method << " " << path << " HTTP/1.1\r\nHost: " << host << ":" << port << "\r\n";
  for (auto& kv : req.headers) {
    if (kv.first == "host" || kv.first == "content-length" || kv.first == "authorization") continue;
    os << kv.first << ": " << kv.second << "\r\n";
  }
  os << "\r\n" << req.body;
  const std::string head = os.str();
  if (send_all(fd, head.data(), head.size())) {
    pollfd fds[2] = {{client, POLLIN, 0}, {fd, POLLIN, 0}};
    char buf[16384];
    bool open = true;
    while (open) {
      const int ready = ::poll(fds, 2, idle_s * 1000);
      if (ready <= 0) break;
      for (int i = 0; i < 2 && open; ++i) {
        if (fds[i].revents & POLLIN) {
          const ssize_t n = ::recv(fds[i].fd, buf, sizeof buf, 0);
          if (n <= 0 || !send_all(fds[1 - i].fd, buf, static_cast<size_t>(n))) open = false;
        } else if (fds[i].revents & (POLLHUP | POLLERR | POLLNVAL)) {
          open = false;
        }
      }
    }
  }
  ::close(fd);
}

// Raw TCP tunnel (`det tunnel`, the ProxyCommand of `det shell`): the client asks with `Upgrade: det-tcp`; the master
// connects to the task's port, answers 101 itself and then copies bytes both ways — nothing of the HTTP request reaches
// the task, which may speak any protocol (sshd). (reference: master/internal/proxy/tcp.go, cli/determined_cli/tunnel.py)
inline void tcp_tunnel(const std::string& host, int port, int client, int idle_s = 3600) {
  auto send_all = [](int fd, const char* p, size_t n) {
    while (n > 0) {
      ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
      if (k <= 0) return false;
      p += k;
      n -= static_cast<size_t>(k);
    }
    return true;
  };
  int fd = ::socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  addr.sin_addr.s_addr = resolve_ipv4(host);
  if (fd < 0 || ::connect(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) < 0) {
    static const char bad[] = "HTTP/1.1 502 Bad Gateway\r\nContent-Length: 0\r\nConnection: close\r\n\r\n";
    send_all(client, bad, sizeof bad - 1);
    if (fd >= 0) ::close(fd);
    return;
  }
  static const char ok[] = "HTTP/1.1 101 Switching Protocols\r\nUpgrade: det-tcp\r\nConnection: Upgrade\r\n\r\n";
  if (send_all(client, ok, sizeof ok - 1)) {
    // each direction ends on its own: an EOF from one side is passed on as a half-close (shutdown of the write side of the
    // other socket) and the opposite direction keeps flowing until it ends too — `echo request | det tunnel` must still
    // receive the answer after its stdin closed
    pollfd fds[2] = {{client, POLLIN, 0}, {fd, POLLIN, 0}};
    bool live[2] = {true, true};
    char buf[16384];
    while (live[0] || live[1]) {
      fds[0].events = live[0] ? POLLIN : 0;
      fds[1].events = live[1] ? POLLIN : 0;
      if (::poll(fds, 2, idle_s * 1000) <= 0) break;
      bool broken = false;
      for (int i = 0; i < 2; ++i) {
        if (!live[i]) continue;
        if (fds[i].revents & (POLLIN | POLLHUP)) {
          const ssize_t n = ::recv(fds[i].fd, buf, sizeof buf, 0);
          if (n > 0) { if (!send_all(fds[1 - i].fd, buf, static_cast<size_t>(n))) broken = true; }
          else { live[i] = false; ::shutdown(fds[1 - i].fd, SHUT_WR); }
        } else if (fds[i].revents & (POLLERR | POLLNVAL)) {
          broken = true;
        }
      }
      if (broken) break;
    }
  }
  ::close(fd);
}

}  // namespace det
