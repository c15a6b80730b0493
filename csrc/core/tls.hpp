// tls.hpp — TLS for the control plane (OpenSSL): the master's HTTPS listener and the clients inside the native binaries
// (agent -> master, master -> Kubernetes API server).
//
// Reference behaviour being matched: the master serves TLS when `security.tls.cert` / `security.tls.key` are configured
// (master/internal/config.go:252-256); agents and harnesses are told to use TLS and get the master's certificate to pin
// (agent/internal/options.go:12-103: security.tls.{enabled, skip_verify, master_cert, master_cert_name}; harness
// DET_USE_TLS / DET_MASTER_CERT_FILE / DET_MASTER_CERT_NAME, common/determined_common/api/request.py:93-215).
//
// A `Stream` is a connected socket that is either plain or wrapped in an SSL session; the HTTP code reads and writes
// through it and never touches OpenSSL itself.
#pragma once
#include <openssl/err.h>
#include <openssl/ssl.h>
#include <openssl/x509v3.h>
#include <sys/socket.h>
#include <unistd.h>

#include <arpa/inet.h>
#include <map>
#include <mutex>
#include <string>

namespace det {

struct TlsClientOptions {
  bool enabled = false;
  bool skip_verify = false;       // encrypt, but accept any certificate
  std::string ca_file;            // certificate (or bundle) the server's chain must lead to: the pinned master certificate
  std::string cert_name;          // name to expect in the certificate instead of the host being dialled
};

inline std::string tls_last_error() {
  char buf[256] = {0};
  const unsigned long e = ERR_get_error();
  if (e == 0) return "no OpenSSL error queued";
  ERR_error_string_n(e, buf, sizeof buf);
  return buf;
}

class Stream {
 public:
  Stream() = default;
  explicit Stream(int fd) : fd_(fd) {}
  Stream(const Stream&) = delete;
  Stream& operator=(const Stream&) = delete;
  Stream(Stream&& o) noexcept : fd_(o.fd_), ssl_(o.ssl_) { o.fd_ = -1; o.ssl_ = nullptr; }
  ~Stream() { close(); }

  int fd() const { return fd_; }
  bool tls() const { return ssl_ != nullptr; }

  // Server side of the handshake on an accepted socket.
  bool accept_tls(SSL_CTX* ctx) {
    ssl_ = SSL_new(ctx);
    if (!ssl_) return false;
    SSL_set_fd(ssl_, fd_);
    if (SSL_accept(ssl_) != 1) { SSL_free(ssl_); ssl_ = nullptr; return false; }
    return true;
  }
  // Client side; verifies against opt.ca_file (or the system store) and the expected name unless skip_verify.
  bool connect_tls(SSL_CTX* ctx, const std::string& host, const TlsClientOptions& opt, std::string* err) {
    ssl_ = SSL_new(ctx);
    if (!ssl_) { if (err) *err = tls_last_error(); return false; }
    SSL_set_fd(ssl_, fd_);
    const std::string name = opt.cert_name.empty() ? host : opt.cert_name;
    SSL_set_tlsext_host_name(ssl_, name.c_str());
    if (!opt.skip_verify) {
      X509_VERIFY_PARAM* vp = SSL_get0_param(ssl_);
      // a dotted quad is matched against IP subjectAltNames, anything else against DNS names
      unsigned char ipbuf[16];
      if (inet_pton(AF_INET, name.c_str(), ipbuf) == 1) X509_VERIFY_PARAM_set1_ip_asc(vp, name.c_str());
      else X509_VERIFY_PARAM_set1_host(vp, name.c_str(), 0);
      SSL_set_verify(ssl_, SSL_VERIFY_PEER, nullptr);
    }
    if (SSL_connect(ssl_) != 1) {
      if (err) {
        const long vr = SSL_get_verify_result(ssl_);
        *err = vr != X509_V_OK ? std::string("certificate verification failed: ") + X509_verify_cert_error_string(vr) : tls_last_error();
      }
      SSL_free(ssl_);
      ssl_ = nullptr;
      return false;
    }
    return true;
  }

  ssize_t read(void* buf, size_t n) {
    if (ssl_) { const int k = SSL_read(ssl_, buf, static_cast<int>(n)); return k; }
    return ::recv(fd_, buf, n, 0);
  }
  bool write_all(const char* p, size_t n) {
    while (n > 0) {
      ssize_t k = ssl_ ? SSL_write(ssl_, p, static_cast<int>(n)) : ::send(fd_, p, n, MSG_NOSIGNAL);
      if (k <= 0) return false;
      p += k;
      n -= static_cast<size_t>(k);
    }
    return true;
  }
  bool pending() const { return ssl_ != nullptr && SSL_pending(ssl_) > 0; }
  void close() {
    if (ssl_) { SSL_shutdown(ssl_); SSL_free(ssl_); ssl_ = nullptr; }
    if (fd_ >= 0) { ::close(fd_); fd_ = -1; }
  }

 private:
  int fd_ = -1;
  SSL* ssl_ = nullptr;
};

// Server context from a PEM certificate chain + private key. Returns nullptr and fills *err on failure.
inline SSL_CTX* tls_server_context(const std::string& cert_file, const std::string& key_file, std::string* err) {
  SSL_CTX* ctx = SSL_CTX_new(TLS_server_method());
  if (!ctx) { if (err) *err = tls_last_error(); return nullptr; }
  SSL_CTX_set_min_proto_version(ctx, TLS1_2_VERSION);
  if (SSL_CTX_use_certificate_chain_file(ctx, cert_file.c_str()) != 1 || SSL_CTX_use_PrivateKey_file(ctx, key_file.c_str(), SSL_FILETYPE_PEM) != 1 ||
      SSL_CTX_check_private_key(ctx) != 1) {
    if (err) *err = "cannot load TLS certificate / key (" + cert_file + ", " + key_file + "): " + tls_last_error();
    SSL_CTX_free(ctx);
    return nullptr;
  }
  return ctx;
}

// One client context per (ca_file) for the life of the process.
inline SSL_CTX* tls_client_context(const TlsClientOptions& opt, std::string* err) {
  static std::mutex mu;
  static std::map<std::string, SSL_CTX*> cache;
  std::lock_guard<std::mutex> g(mu);
  auto it = cache.find(opt.ca_file);
  if (it != cache.end()) return it->second;
  SSL_CTX* ctx = SSL_CTX_new(TLS_client_method());
  if (!ctx) { if (err) *err = tls_last_error(); return nullptr; }
  SSL_CTX_set_min_proto_version(ctx, TLS1_2_VERSION);
  if (!opt.ca_file.empty()) {
    if (SSL_CTX_load_verify_locations(ctx, opt.ca_file.c_str(), nullptr) != 1) {
      if (err) *err = "cannot read certificate file " + opt.ca_file + ": " + tls_last_error();
      SSL_CTX_free(ctx);
      return nullptr;
    }
  } else {
    SSL_CTX_set_default_verify_paths(ctx);
  }
  cache[opt.ca_file] = ctx;
  return ctx;
}

}  // namespace det
