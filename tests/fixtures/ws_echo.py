"""A service with one upgradeable endpoint: plain requests get "ok"; a request carrying `Upgrade:` gets `101 Switching
Protocols` and the connection then echoes every chunk upper-cased (a stand-in for a WebSocket application: the proxy under
test copies bytes and does not interpret frames)."""
import socket
import sys
import threading


def serve(conn):
    buf = b""
    while b"\r\n\r\n" not in buf:
        d = conn.recv(4096)
        if not d:
            conn.close()
            return
        buf += d
    head = buf.split(b"\r\n\r\n")[0].decode("latin1")
    lines = head.split("\r\n")
    hdrs = {l.split(":", 1)[0].strip().lower(): l.split(":", 1)[1].strip() for l in lines[1:] if ":" in l}
    if "upgrade" in hdrs:
        path = lines[0].split()[1]
        conn.sendall(("HTTP/1.1 101 Switching Protocols\r\nUpgrade: %s\r\nConnection: Upgrade\r\nX-Path: %s\r\nX-Key: %s\r\n\r\n"
                      % (hdrs["upgrade"], path, hdrs.get("sec-websocket-key", ""))).encode())
        while True:
            d = conn.recv(4096)
            if not d or d.strip() == b"bye":
                break
            conn.sendall(d.upper())
    else:
        conn.sendall(b"HTTP/1.1 200 OK\r\nContent-Type: text/plain\r\nContent-Length: 2\r\nConnection: close\r\n\r\nok")
    conn.close()


srv = socket.socket()
srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
srv.bind(("127.0.0.1", int(sys.argv[1])))
srv.listen(16)
while True:
    c, _ = srv.accept()
    threading.Thread(target=serve, args=(c,), daemon=True).start()
