// qtshim.h — TEST INFRASTRUCTURE. Minimal stand-ins for the Qt classes the reference's render path touches,
// so that its OWN sources (rasterizer.cpp, galaxycomponent(s).cpp, galaxy.cpp, gamercamera.cpp, buffer2d.cpp,
// noise.cpp ...) compile in place into oracle/_ref/ without Qt (oracle/Makefile, target `ref`).
//
// What this is and is not: the CONTROL FLOW AND FORMULAS that end up in oracle/_ref/libgamer_ref_path.so are the
// reference's own, byte for byte; the Qt VALUE-CLASS ARITHMETIC underneath (QVector3D, QQuaternion, QMatrix4x4)
// is the Qt 5.15 restatement of oracle/qt5_math.h, the same one the oracle uses. Comparing the oracle with
// that library therefore pins the oracle's reading of the reference's code, not Qt's arithmetic.
// Containers, strings, files and threads are thin wrappers over the C++ standard library; GUI classes are
// empty. Nothing here is used by the product.
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../qt5_math.h"
using q5::M4;
using q5::Quat;
using q5::V3;

typedef int32_t qint32;
typedef uint32_t quint32;
typedef int64_t qint64;
typedef uint64_t quint64;
typedef double qreal;
typedef unsigned int QRgb;
typedef unsigned char uchar;

#define Q_OBJECT
#define Q_ENUMS(x)
#define Q_UNUSED(x) (void)x;
#define slots
#define signals public
#define emit

// ---- QVector / QList ------------------------------------------------------------------------------------
template <class T>
class QVector {
    std::vector<T> v;

public:
    QVector() {}
    QVector(int n) : v(size_t(n)) {}
    QVector(int n, const T &x) : v(size_t(n), x) {}
    QVector(std::initializer_list<T> l) : v(l) {}
    int size() const { return int(v.size()); }
    int count() const { return int(v.size()); }
    int length() const { return int(v.size()); }
    bool isEmpty() const { return v.empty(); }
    void resize(int n) { v.resize(size_t(n)); }
    void reserve(int n) { v.reserve(size_t(n)); }
    void clear() { v.clear(); }
    void append(const T &x) { v.push_back(x); }
    void push_back(const T &x) { v.push_back(x); }
    void push_front(const T &x) { v.insert(v.begin(), x); }
    void prepend(const T &x) { v.insert(v.begin(), x); }
    void insert(int i, const T &x) { v.insert(v.begin() + i, x); }
    void remove(int i) { v.erase(v.begin() + i); }
    void removeAt(int i) { v.erase(v.begin() + i); }
    void removeFirst() { v.erase(v.begin()); }
    void removeLast() { v.pop_back(); }
    int removeAll(const T &x) {
        auto e = std::remove(v.begin(), v.end(), x);
        int n = int(v.end() - e);
        v.erase(e, v.end());
        return n;
    }
    bool contains(const T &x) const { return std::find(v.begin(), v.end(), x) != v.end(); }
    int indexOf(const T &x) const {
        auto it = std::find(v.begin(), v.end(), x);
        return it == v.end() ? -1 : int(it - v.begin());
    }
    T &operator[](int i) { return v[size_t(i)]; }
    const T &operator[](int i) const { return v[size_t(i)]; }
    const T &at(int i) const { return v[size_t(i)]; }
    T value(int i) const { return (i >= 0 && i < size()) ? v[size_t(i)] : T(); }
    T &first() { return v.front(); }
    T &last() { return v.back(); }
    const T &first() const { return v.front(); }
    const T &last() const { return v.back(); }
    void fill(const T &x) { std::fill(v.begin(), v.end(), x); }
    T *data() { return v.data(); }
    const T *data() const { return v.data(); }
    typename std::vector<T>::iterator begin() { return v.begin(); }
    typename std::vector<T>::iterator end() { return v.end(); }
    typename std::vector<T>::const_iterator begin() const { return v.begin(); }
    typename std::vector<T>::const_iterator end() const { return v.end(); }
    QVector<T> &operator<<(const T &x) { v.push_back(x); return *this; }
    bool operator==(const QVector<T> &o) const { return v == o.v; }
};
template <class T> using QList = QVector<T>;

// ---- QString (UTF-8 inside) / QStringList / QByteArray ---------------------------------------------------
class QString;
typedef QVector<QString> QStringList;
class QString {
    std::string s;

public:
    QString() {}
    QString(const char *c) : s(c ? c : "") {}
    QString(const std::string &c) : s(c) {}
    static QString fromStdString(const std::string &c) { return QString(c); }
    static QString number(int v) { return QString(std::to_string(v)); }
    QString rightJustified(int w, char fill) const { return int(s.size()) >= w ? *this : QString(std::string(size_t(w) - s.size(), fill) + s); }
    static QString number(unsigned v) { return QString(std::to_string(v)); }
    static QString number(long v) { return QString(std::to_string(v)); }
    static QString number(long long v) { return QString(std::to_string(v)); }
    static QString number(unsigned long v) { return QString(std::to_string(v)); }
    static QString number(double v) { char b[64]; std::snprintf(b, sizeof b, "%g", v); return QString(b); }
    std::string toStdString() const { return s; }
    const std::string &std_ref() const { return s; }
    QString toLower() const { QString r(*this); for (char &c : r.s) c = char(std::tolower((unsigned char)c)); return r; }
    QString toUpper() const { QString r(*this); for (char &c : r.s) c = char(std::toupper((unsigned char)c)); return r; }
    QString trimmed() const {
        size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
        return a == std::string::npos ? QString() : QString(s.substr(a, b - a + 1));
    }
    float toFloat(bool *ok = nullptr) const { if (ok) *ok = true; return std::strtof(s.c_str(), nullptr); }
    double toDouble(bool *ok = nullptr) const { if (ok) *ok = true; return std::strtod(s.c_str(), nullptr); }
    int toInt(bool *ok = nullptr) const { if (ok) *ok = true; return int(std::strtol(s.c_str(), nullptr, 10)); }
    bool isEmpty() const { return s.empty(); }
    bool isNull() const { return s.empty(); }
    int size() const { return int(s.size()); }
    int length() const { return int(s.size()); }
    bool contains(const QString &o) const { return s.find(o.s) != std::string::npos; }
    bool startsWith(const QString &o) const { return s.rfind(o.s, 0) == 0; }
    bool endsWith(const QString &o) const { return s.size() >= o.s.size() && s.compare(s.size() - o.s.size(), o.s.size(), o.s) == 0; }
    QString &replace(const QString &a, const QString &b) {
        if (a.s.empty()) return *this;
        for (size_t p = 0; (p = s.find(a.s, p)) != std::string::npos; p += b.s.size()) s.replace(p, a.s.size(), b.s);
        return *this;
    }
    QString remove(const QString &a) const { QString r(*this); r.replace(a, QString()); return r; }
    QStringList split(const QString &sep) const;
    QString &operator+=(const QString &o) { s += o.s; return *this; }
    bool operator==(const QString &o) const { return s == o.s; }
    bool operator!=(const QString &o) const { return s != o.s; }
    bool operator<(const QString &o) const { return s < o.s; }
    friend QString operator+(const QString &a, const QString &b) { return QString(a.s + b.s); }
    friend QString operator+(const char *a, const QString &b) { return QString(std::string(a) + b.s); }
    friend QString operator+(const QString &a, const char *b) { return QString(a.s + b); }
};
inline QStringList QString::split(const QString &sep) const {
    QStringList out;
    size_t p = 0, q;
    while ((q = s.find(sep.s, p)) != std::string::npos) { out.append(QString(s.substr(p, q - p))); p = q + sep.s.size(); }
    out.append(QString(s.substr(p)));
    return out;
}

class QByteArray {
    std::vector<char> d;

public:
    QByteArray() {}
    QByteArray &append(const char *p, int n) { d.insert(d.end(), p, p + n); return *this; }
    QByteArray &append(char c) { d.push_back(c); return *this; }
    int size() const { return int(d.size()); }
    char *data() { return d.data(); }
    const char *data() const { return d.data(); }
    void clear() { d.clear(); }
};

// ---- QFile / QDataStream (big-endian, floats widened to doubles: Qt's defaults) -----------------------------
struct QIODevice { enum OpenModeFlag { ReadOnly = 1, WriteOnly = 2, ReadWrite = 3, Text = 16 }; };
class QFile {
    std::string name;
    FILE *f = nullptr;

public:
    QFile() {}
    QFile(const QString &n) : name(n.toStdString()) {}
    ~QFile() { close(); }
    void setFileName(const QString &n) { name = n.toStdString(); }
    QString fileName() const { return QString(name); }
    bool open(int mode) { close(); f = std::fopen(name.c_str(), (mode & QIODevice::WriteOnly) ? "wb" : "rb"); return f != nullptr; }
    bool isOpen() const { return f != nullptr; }
    void close() { if (f) std::fclose(f); f = nullptr; }
    void flush() { if (f) std::fflush(f); }
    bool exists() const { FILE *t = std::fopen(name.c_str(), "rb"); if (t) std::fclose(t); return t != nullptr; }
    static bool exists(const QString &n) { return QFile(n).exists(); }
    bool atEnd() const { if (!f) return true; int c = std::fgetc(f); if (c == EOF) return true; std::ungetc(c, f); return false; }
    FILE *handle() { return f; }
};

class QDataStream {
    FILE *f = nullptr;
    QByteArray *ba = nullptr;
    bool bad = false;
    void put(const void *p, size_t n) { if (f) std::fwrite(p, 1, n, f); else if (ba) ba->append((const char *)p, int(n)); }
    void get(void *p, size_t n) { if (!f || std::fread(p, 1, n, f) != n) { std::memset(p, 0, n); bad = true; } }
    void put_be(uint64_t v, int n) { unsigned char b[8]; for (int i = 0; i < n; i++) b[i] = (unsigned char)(v >> (8 * (n - 1 - i))); put(b, size_t(n)); }
    uint64_t get_be(int n) { unsigned char b[8]; get(b, size_t(n)); uint64_t v = 0; for (int i = 0; i < n; i++) v = (v << 8) | b[i]; return v; }

public:
    enum Version { Qt_4_0 = 7, Qt_5_0 = 13, Qt_5_6 = 17, Qt_5_12 = 18, Qt_5_15 = 19, Qt_6_0 = 20 };
    enum Status { Ok, ReadPastEnd };
    QDataStream() {}
    QDataStream(QFile *file) : f(file->handle()) {}
    QDataStream(QByteArray *b, int) : ba(b) {}
    void setVersion(int) {}
    Status status() const { return bad ? ReadPastEnd : Ok; }
    bool atEnd() const { if (!f) return true; int c = std::fgetc(f); if (c == EOF) return true; std::ungetc(c, f); return false; }
    int writeRawData(const char *p, int n) { put(p, size_t(n)); return n; }
    int readRawData(char *p, int n) { get(p, size_t(n)); return n; }
    QDataStream &operator<<(qint32 v) { put_be(uint32_t(v), 4); return *this; }
    QDataStream &operator<<(quint32 v) { put_be(v, 4); return *this; }
    QDataStream &operator<<(bool v) { put_be(v ? 1 : 0, 1); return *this; }
    QDataStream &operator<<(double v) { uint64_t b; std::memcpy(&b, &v, 8); put_be(b, 8); return *this; }
    QDataStream &operator<<(float v) { return *this << double(v); }
    QDataStream &operator<<(const QString &s) {   // UTF-16BE; the shipped files are ASCII
        const std::string &u = s.std_ref();
        put_be(uint32_t(u.size() * 2), 4);
        for (unsigned char c : u) put_be(c, 2);
        return *this;
    }
    QDataStream &operator>>(qint32 &v) { v = qint32(uint32_t(get_be(4))); return *this; }
    QDataStream &operator>>(quint32 &v) { v = quint32(get_be(4)); return *this; }
    QDataStream &operator>>(bool &v) { v = get_be(1) != 0; return *this; }
    QDataStream &operator>>(double &v) { uint64_t b = get_be(8); std::memcpy(&v, &b, 8); return *this; }
    QDataStream &operator>>(float &v) { double d; *this >> d; v = float(d); return *this; }
    QDataStream &operator>>(QString &s) {
        uint32_t n = uint32_t(get_be(4));
        std::string u;
        if (n != 0xFFFFFFFFu && !bad)
            for (uint32_t i = 0; i + 1 < n && !bad; i += 2) u += char(get_be(2) & 0xFF);
        s = QString(u);
        return *this;
    }
    template <class T> QDataStream &operator<<(const QVector<T> &v) {
        *this << quint32(v.size());
        for (int i = 0; i < v.size(); i++) { T t = v[i]; *this << t; }
        return *this;
    }
    template <class T> QDataStream &operator>>(QVector<T> &v) {
        quint32 n = 0;
        *this >> n;
        v.clear();
        for (quint32 i = 0; i < n && !bad; i++) { T t; *this >> t; v.append(t); }
        return *this;
    }
};

// ---- value classes: Qt's API over the Qt 5.15 arithmetic of oracle/qt5_math.h ------------------------------
class QVector3D {
public:
    V3 v;
    QVector3D() {}
    QVector3D(float x, float y, float z) : v(x, y, z) {}
    QVector3D(const V3 &a) : v(a) {}
    float x() const { return v.x; }
    float y() const { return v.y; }
    float z() const { return v.z; }
    void setX(float a) { v.x = a; }
    void setY(float a) { v.y = a; }
    void setZ(float a) { v.z = a; }
    float &operator[](int i) { return i == 0 ? v.x : i == 1 ? v.y : v.z; }
    float operator[](int i) const { return i == 0 ? v.x : i == 1 ? v.y : v.z; }
    float length() const { return q5::length(v); }
    float lengthSquared() const { return v.x * v.x + v.y * v.y + v.z * v.z; }
    QVector3D normalized() const { return QVector3D(q5::normalized(v)); }
    void normalize() { v = q5::normalized(v); }
    bool isNull() const { return v.x == 0.f && v.y == 0.f && v.z == 0.f; }
    static float dotProduct(const QVector3D &a, const QVector3D &b) { return q5::dot(a.v, b.v); }
    static QVector3D crossProduct(const QVector3D &a, const QVector3D &b) { return QVector3D(q5::cross(a.v, b.v)); }
    float distanceToPoint(const QVector3D &o) const { return q5::length(v - o.v); }
    QVector3D &operator+=(const QVector3D &o) { v = v + o.v; return *this; }
    QVector3D &operator-=(const QVector3D &o) { v = v - o.v; return *this; }
    QVector3D &operator*=(float f) { v = v * f; return *this; }
    QVector3D &operator*=(const QVector3D &o) { v = V3(v.x * o.v.x, v.y * o.v.y, v.z * o.v.z); return *this; }
    QVector3D &operator/=(float f) { v = v / f; return *this; }
    friend QVector3D operator+(const QVector3D &a, const QVector3D &b) { return QVector3D(a.v + b.v); }
    friend QVector3D operator-(const QVector3D &a, const QVector3D &b) { return QVector3D(a.v - b.v); }
    friend QVector3D operator-(const QVector3D &a) { return QVector3D(-a.v); }
    friend QVector3D operator*(const QVector3D &a, float f) { return QVector3D(a.v * f); }
    friend QVector3D operator*(float f, const QVector3D &a) { return QVector3D(f * a.v); }
    friend QVector3D operator*(const QVector3D &a, const QVector3D &b) { return QVector3D(a.v.x * b.v.x, a.v.y * b.v.y, a.v.z * b.v.z); }
    friend QVector3D operator/(const QVector3D &a, float f) { return QVector3D(a.v / f); }
    friend bool operator==(const QVector3D &a, const QVector3D &b) { return a.v.x == b.v.x && a.v.y == b.v.y && a.v.z == b.v.z; }
    friend bool operator!=(const QVector3D &a, const QVector3D &b) { return !(a == b); }
};
inline QDataStream &operator<<(QDataStream &s, const QVector3D &v) { return s << v.x() << v.y() << v.z(); }
inline QDataStream &operator>>(QDataStream &s, QVector3D &v) { float x, y, z; s >> x >> y >> z; v = QVector3D(x, y, z); return s; }

class QVector2D {
public:
    float a = 0, b = 0;
    QVector2D() {}
    QVector2D(float x, float y) : a(x), b(y) {}
    float x() const { return a; }
    float y() const { return b; }
};

class QVector4D {
public:
    float c[4] = {0, 0, 0, 0};
    QVector4D() {}
    QVector4D(float x, float y, float z, float w) : c{x, y, z, w} {}
    QVector4D(const QVector3D &v, float w) : c{v.x(), v.y(), v.z(), w} {}
    float x() const { return c[0]; }
    float y() const { return c[1]; }
    float z() const { return c[2]; }
    float w() const { return c[3]; }
    QVector3D toVector3D() const { return QVector3D(c[0], c[1], c[2]); }
    QVector3D toVector3DAffine() const { return QVector3D(c[0] / c[3], c[1] / c[3], c[2] / c[3]); }
};

class QQuaternion {
public:
    Quat q;
    QQuaternion() {}
    QQuaternion(float w, float x, float y, float z) : q(w, x, y, z) {}
    QQuaternion(const Quat &a) : q(a) {}
    float scalar() const { return q.w; }
    float x() const { return q.x; }
    float y() const { return q.y; }
    float z() const { return q.z; }
    QQuaternion normalized() const { return QQuaternion(q5::normalized(q)); }
    QQuaternion conjugated() const { return QQuaternion(q.w, -q.x, -q.y, -q.z); }
    QVector3D rotatedVector(const QVector3D &v) const { return QVector3D(q5::rotate(q, v.v)); }
    static QQuaternion fromAxisAndAngle(const QVector3D &axis, float deg) { return QQuaternion(q5::from_axis_angle(axis.v, deg)); }
    static QQuaternion fromAxisAndAngle(float x, float y, float z, float deg) { return fromAxisAndAngle(QVector3D(x, y, z), deg); }
    static QQuaternion rotationTo(const QVector3D &from, const QVector3D &to) { return QQuaternion(q5::rotation_to(from.v, to.v)); }
    static QQuaternion fromEulerAngles(float, float, float) { return QQuaternion(); }     // GUI camera only, off the path
    static QQuaternion fromEulerAngles(const QVector3D &) { return QQuaternion(); }
    class QMatrix3x3 toRotationMatrix() const;
    friend QQuaternion operator*(const QQuaternion &a, const QQuaternion &b) { return QQuaternion(q5::mul(a.q, b.q)); }
    friend QVector3D operator*(const QQuaternion &a, const QVector3D &v) { return a.rotatedVector(v); }
};

class QMatrix3x3 {
public:
    float m[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    float &operator()(int r, int c) { return m[c * 3 + r]; }
    float operator()(int r, int c) const { return m[c * 3 + r]; }
    void setToIdentity() { for (int i = 0; i < 9; i++) m[i] = (i % 4 == 0) ? 1.f : 0.f; }
};

inline QMatrix3x3 QQuaternion::toRotationMatrix() const { return QMatrix3x3(); }               // GUI camera only, off the path

class QMatrix4x4 {
public:
    M4 a;
    QMatrix4x4() : a(M4::identity()) {}
    QMatrix4x4(const M4 &m) : a(m) {}
    void setToIdentity() { a = M4::identity(); }
    float &operator()(int r, int c) { return a.m[c][r]; }
    float operator()(int r, int c) const { return a.m[c][r]; }
    void perspective(float vdeg, float aspect, float zn, float zf) { a = q5::mul(a, q5::perspective(vdeg, aspect, zn, zf)); }
    void lookAt(const QVector3D &eye, const QVector3D &center, const QVector3D &up) { a = q5::mul(a, q5::look_at(eye.v, center.v, up.v)); }
    QMatrix4x4 inverted(bool *ok = nullptr) const { if (ok) *ok = true; return QMatrix4x4(q5::inverted(a)); }
    QMatrix4x4 transposed() const { QMatrix4x4 r; for (int c = 0; c < 4; c++) for (int w = 0; w < 4; w++) r.a.m[c][w] = a.m[w][c]; return r; }
    void rotate(float, float, float, float) {}                 // GUI camera only, off the path
    void rotate(float, const QVector3D &) {}
    void rotate(const QQuaternion &) {}
    void translate(const QVector3D &) {}
    void scale(float) {}
    QMatrix3x3 normalMatrix() const { return QMatrix3x3(); }
    QVector4D column(int c) const { return QVector4D(a.m[c][0], a.m[c][1], a.m[c][2], a.m[c][3]); }
    QVector4D row(int r) const { return QVector4D(a.m[0][r], a.m[1][r], a.m[2][r], a.m[3][r]); }
    friend QMatrix4x4 operator*(const QMatrix4x4 &x, const QMatrix4x4 &y) { return QMatrix4x4(q5::mul(x.a, y.a)); }
    friend QVector4D operator*(const QMatrix4x4 &m, const QVector4D &v) {   // Qt: row sums left to right, float
        QVector4D r;
        for (int w = 0; w < 4; w++) r.c[w] = v.c[0] * m.a.m[0][w] + v.c[1] * m.a.m[1][w] + v.c[2] * m.a.m[2][w] + v.c[3] * m.a.m[3][w];
        return r;
    }
    friend QVector3D operator*(const QMatrix4x4 &m, const QVector3D &v) { return (m * QVector4D(v, 1.0f)).toVector3DAffine(); }
};

// ---- colours and images -------------------------------------------------------------------------------------
inline QRgb qRgba(int r, int g, int b, int a) { return ((a & 0xffu) << 24) | ((r & 0xffu) << 16) | ((g & 0xffu) << 8) | (b & 0xffu); }
inline QRgb qRgb(int r, int g, int b) { return qRgba(r, g, b, 255); }
inline int qRed(QRgb c) { return (c >> 16) & 0xff; }
inline int qGreen(QRgb c) { return (c >> 8) & 0xff; }
inline int qBlue(QRgb c) { return c & 0xff; }
inline int qAlpha(QRgb c) { return c >> 24; }
class QColor {
    int r_ = 0, g_ = 0, b_ = 0, a_ = 255;

public:
    QColor() {}
    QColor(int r, int g, int b, int a = 255) : r_(r), g_(g), b_(b), a_(a) {}
    QColor(QRgb c) : r_(qRed(c)), g_(qGreen(c)), b_(qBlue(c)), a_(255) {}
    static QColor fromRgba(QRgb c) { return QColor(qRed(c), qGreen(c), qBlue(c), qAlpha(c)); }
    static QColor fromRgb(int r, int g, int b) { return QColor(r, g, b); }
    int red() const { return r_; }
    int green() const { return g_; }
    int blue() const { return b_; }
    int alpha() const { return a_; }
    QRgb rgb() const { return qRgb(r_, g_, b_); }
    QRgb rgba() const { return qRgba(r_, g_, b_, a_); }
};
class QImage {
    int w_ = 0, h_ = 0;
    std::vector<QRgb> px;

public:
    enum Format { Format_Invalid, Format_RGB32, Format_ARGB32 };
    QImage() {}
    QImage(int w, int h, Format) : w_(w), h_(h), px(size_t(w) * size_t(h), 0xff000000u) {}
    int width() const { return w_; }
    int height() const { return h_; }
    void fill(const QColor &c) { std::fill(px.begin(), px.end(), c.rgba()); }
    void fill(unsigned c) { std::fill(px.begin(), px.end(), QRgb(c)); }
    void setPixel(int x, int y, QRgb c) { if (x >= 0 && x < w_ && y >= 0 && y < h_) px[size_t(y) * w_ + x] = c; }
    void setPixelColor(int x, int y, const QColor &c) { setPixel(x, y, c.rgba()); }
    QRgb pixel(int x, int y) const { return px[size_t(y) * w_ + x]; }
    QColor pixelColor(int x, int y) const { return QColor::fromRgba(pixel(x, y)); }
    uchar *bits() { return reinterpret_cast<uchar *>(px.data()); }
    const uchar *bits() const { return reinterpret_cast<const uchar *>(px.data()); }
    uchar *scanLine(int y) { return bits() + size_t(y) * w_ * 4; }
    bool save(const QString &) const { return true; }
    bool load(const QString &) { return false; }
    QImage scaled(int, int) const { return *this; }
    QImage copy() const { return *this; }
    bool isNull() const { return px.empty(); }
};

// ---- threads, timers, messages: inert ------------------------------------------------------------------------
class QObject {
public:
    virtual ~QObject() {}
};
class QThread : public QObject {
    bool interrupt_ = false;

public:
    enum Priority { IdlePriority, LowestPriority, LowPriority, NormalPriority, HighPriority, HighestPriority, TimeCriticalPriority, InheritPriority };
    virtual ~QThread() {}
    virtual void run() {}
    void start(Priority = InheritPriority) { interrupt_ = false; run(); }    // synchronous: the caller wants the pixels
    bool isRunning() const { return false; }
    bool isFinished() const { return true; }
    bool wait(unsigned long = 0) { return true; }
    void quit() {}
    void exit(int = 0) {}
    void terminate() {}
    void requestInterruption() { interrupt_ = true; }
    bool isInterruptionRequested() const { return interrupt_; }
    static void msleep(unsigned long) {}
    static void sleep(unsigned long) {}
    static QThread *currentThread() { static QThread t; return &t; }
};
class QMutex {
public:
    void lock() {}
    void unlock() {}
    bool tryLock() { return true; }
};
class QMutexLocker {
public:
    QMutexLocker(QMutex *) {}
};
class QWaitCondition {
public:
    void wakeOne() {}
    void wakeAll() {}
    bool wait(QMutex *, unsigned long = 0) { return true; }
};
class QElapsedTimer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();

public:
    void start() { t0 = std::chrono::steady_clock::now(); }
    qint64 restart() { qint64 e = elapsed(); start(); return e; }
    qint64 elapsed() const { return std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count(); }
    qint64 nsecsElapsed() const { return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count(); }
    bool isValid() const { return true; }
};
class QDateTime {
public:
    static QDateTime currentDateTime() { return QDateTime(); }
    QString toString() const { return QString(""); }
    QString toString(const QString &) const { return QString(""); }
};
class QListWidgetItem {
public:
    void setForeground(const QColor &) {}
};
class QListWidget {
    QListWidgetItem it;

public:
    void clear() {}
    void addItem(const QString &) {}
    QListWidgetItem *item(int) { return &it; }
    void scrollToBottom() {}
};
struct QDir { enum Filter { Files = 2, Dirs = 1, NoDotAndDotDot = 0x1000 }; };
class QDirIterator {
public:
    enum IteratorFlag { NoIteratorFlags = 0, Subdirectories = 2 };
    QDirIterator(const QString &, int = 0, int = 0) {}
    QDirIterator(const QString &, const QStringList &, int = 0, int = 0) {}
    bool hasNext() const { return false; }
    QString next() { return QString(); }
};
class QTextStream {
public:
    QTextStream() {}
    QTextStream(QFile *) {}
    bool atEnd() const { return true; }
    QString readLine() { return QString(); }
    QString readAll() { return QString(); }
    template <class T> QTextStream &operator<<(const T &) { return *this; }
};
class QOpenGLFunctions {};
class QDebug {
public:
    template <class T> QDebug &operator<<(const T &) { return *this; }
    QDebug &noquote() { return *this; }
    QDebug &nospace() { return *this; }
};
inline QDebug qDebug() { return QDebug(); }
inline QDebug qWarning() { return QDebug(); }
inline QDebug qCritical() { return QDebug(); }
template <class T> inline const T &qMin(const T &a, const T &b) { return (a < b) ? a : b; }
template <class T> inline const T &qMax(const T &a, const T &b) { return (a < b) ? b : a; }
template <class T> inline T qAbs(const T &a) { return a < 0 ? -a : a; }
inline float qDegreesToRadians(float d) { return d * float(3.14159265358979323846 / 180.0); }
inline double qDegreesToRadians(double d) { return d * (3.14159265358979323846 / 180.0); }
inline float qRadiansToDegrees(float r) { return r * float(180.0 / 3.14159265358979323846); }
inline double qRadiansToDegrees(double r) { return r * (180.0 / 3.14159265358979323846); }
