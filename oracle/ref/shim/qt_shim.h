// oracle/ref/shim/qt_shim.h — TEST INFRASTRUCTURE.
// Just enough of the Qt API surface, as plain C++, for the reference's OWN hot-path sources (lib/SamplerSynthVoice.cpp,
// lib/SamplerSynth.cpp, lib/JackPassthrough.cpp, lib/ClipAudioSourcePositionsModel.cpp, lib/SyncTimer.cpp) to be compiled
// in place by oracle/Makefile into oracle/_ref/.  Qt5 is not installed here; none of the arithmetic the oracle is pinned
// against lives in Qt - these classes are containers, strings, the object tree and the signal/slot declarations (signals
// have empty bodies in ref/libzl_ref.cpp, where moc would have generated them).
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <initializer_list>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <type_traits>
#include <unordered_map>
#include <vector>

typedef int64_t qint64; typedef uint64_t quint64; typedef int32_t qint32; typedef uint32_t quint32;
typedef uint16_t quint16; typedef int16_t qint16; typedef uint8_t quint8; typedef int8_t qint8; typedef double qreal;
typedef unsigned int uint; typedef unsigned char uchar;

#define Q_OBJECT
#define Q_PROPERTY(...)
#define Q_INVOKABLE
#define Q_SIGNAL
#define Q_SLOT
#define Q_SIGNALS public
#define Q_SLOTS
#define Q_EMIT
#define Q_FUNC_INFO __PRETTY_FUNCTION__
#define Q_UNUSED(x) (void) x;
#define Q_DECLARE_METATYPE(x)
#define Q_ENUM(x)
#define Q_DISABLE_COPY(x)
#define Q_DECL_OVERRIDE override

namespace Qt {
enum ItemDataRole { DisplayRole = 0, UserRole = 0x0100 };
enum ConnectionType { AutoConnection, DirectConnection, QueuedConnection, BlockingQueuedConnection };
enum TimerType { PreciseTimer, CoarseTimer };
enum DateFormat { TextDate, ISODate };
}

template <typename T> inline const T& qMin(const T& a, const T& b) { return (a < b) ? a : b; }
template <typename T> inline const T& qMax(const T& a, const T& b) { return (a < b) ? b : a; }
template <typename T> inline const T& qBound(const T& lo, const T& v, const T& hi) { return qMax(lo, qMin(hi, v)); }
template <typename T> inline typename std::add_const<T>::type& qAsConst(T& t) { return t; }
template <typename C> inline void qDeleteAll(const C& c) { for (auto* p : c) delete p; }

class QByteArray {
public:
    QByteArray() {}
    QByteArray(const char* s) : s_(s ? s : "") {}
    QByteArray(const std::string& s) : s_(s) {}
    const char* constData() const { return s_.c_str(); }
    const char* data() const { return s_.c_str(); }
    operator const char*() const { return s_.c_str(); }
    int size() const { return (int) s_.size(); }
    bool operator==(const QByteArray& o) const { return s_ == o.s_; }
private:
    std::string s_;
};

class QString {
public:
    QString() {}
    QString(const char* s) : s_(s ? s : "") {}
    QString(const QByteArray& b) : s_(b.constData()) {}
    QString(const std::string& s) : s_(s) {}
    QByteArray toUtf8() const { return QByteArray(s_); }
    QByteArray toLatin1() const { return QByteArray(s_); }
    std::string toStdString() const { return s_; }
    bool isEmpty() const { return s_.empty(); }
    int length() const { return (int) s_.size(); }
    QString arg(const QString& a) const {            // replaces the lowest-numbered %N marker
        int best = -1; size_t at = std::string::npos, len = 0;
        for (size_t i = 0; i + 1 < s_.size(); ++i)
            if (s_[i] == '%' && s_[i + 1] >= '1' && s_[i + 1] <= '9') {
                size_t j = i + 1; int n = 0;
                while (j < s_.size() && s_[j] >= '0' && s_[j] <= '9') { n = n * 10 + (s_[j] - '0'); ++j; }
                if (best < 0 || n < best) { best = n; at = i; len = j - i; }
            }
        if (best < 0) return *this;
        std::string marker = s_.substr(at, len), out = s_;
        for (size_t p = out.find(marker); p != std::string::npos; p = out.find(marker, p + a.s_.size())) out.replace(p, marker.size(), a.s_);
        return QString(out);
    }
    QString arg(int v) const { return arg(QString(std::to_string(v))); }
    QString arg(long long v) const { return arg(QString(std::to_string(v))); }
    QString arg(unsigned long long v) const { return arg(QString(std::to_string(v))); }
    QString arg(double v) const { return arg(QString(std::to_string(v))); }
    static QString number(int v) { return QString(std::to_string(v)); }
    static QString number(long long v) { return QString(std::to_string(v)); }
    static QString number(double v) { return QString(std::to_string(v)); }
    static QString fromUtf8(const char* s) { return QString(s); }
    bool operator==(const QString& o) const { return s_ == o.s_; }
    bool operator!=(const QString& o) const { return s_ != o.s_; }
    bool operator<(const QString& o) const { return s_ < o.s_; }
    QString operator+(const QString& o) const { return QString(s_ + o.s_); }
    QString& operator+=(const QString& o) { s_ += o.s_; return *this; }
    bool startsWith(const QString& p) const { return s_.compare(0, p.s_.size(), p.s_) == 0; }
    bool endsWith(const QString& p) const { return s_.size() >= p.s_.size() && s_.compare(s_.size() - p.s_.size(), p.s_.size(), p.s_) == 0; }
    QString left(int n) const { return n < 0 ? *this : QString(s_.substr(0, (size_t) n)); }
    QString mid(int pos, int n = -1) const { return QString(n < 0 ? s_.substr((size_t) pos) : s_.substr((size_t) pos, (size_t) n)); }
    int lastIndexOf(char c) const { const size_t p = s_.rfind(c); return p == std::string::npos ? -1 : (int) p; }
    int indexOf(char c) const { const size_t p = s_.find(c); return p == std::string::npos ? -1 : (int) p; }
    bool contains(const QString& o) const { return s_.find(o.s_) != std::string::npos; }
    const std::string& str() const { return s_; }
private:
    std::string s_;
};
typedef QString QLatin1String;
typedef QString QStringLiteralType;
#define QStringLiteral(x) QString(x)
namespace std { template <> struct hash<QString> { size_t operator()(const QString& s) const { return hash<string>()(s.str()); } }; }

// qDebug() << ... : swallowed
struct QDebugSink {
    template <typename T> QDebugSink& operator<<(const T&) { return *this; }
    QDebugSink& nospace() { return *this; }
    QDebugSink& noquote() { return *this; }
};
inline QDebugSink qDebug() { return QDebugSink(); }
inline QDebugSink qInfo() { return QDebugSink(); }
inline QDebugSink qWarning() { return QDebugSink(); }
inline QDebugSink qCritical() { return QDebugSink(); }

template <typename T> class QList : public std::vector<T> {
public:
    using std::vector<T>::vector;
    QList() {}
    QList(std::initializer_list<T> l) : std::vector<T>(l) {}
    QList& operator<<(const T& v) { this->push_back(v); return *this; }
    void append(const T& v) { this->push_back(v); }
    void prepend(const T& v) { this->insert(this->begin(), v); }
    int count() const { return (int) this->size(); }
    int length() const { return (int) this->size(); }
    bool isEmpty() const { return this->empty(); }
    const T& at(int i) const { return (*this)[(size_t) i]; }
    T value(int i, const T& d = T()) const { return (i >= 0 && i < count()) ? (*this)[(size_t) i] : d; }
    void replace(int i, const T& v) { (*this)[(size_t) i] = v; }
    void removeAt(int i) { this->erase(this->begin() + i); }
    T takeFirst() { T v = this->front(); this->erase(this->begin()); return v; }
    T takeAt(int i) { T v = (*this)[(size_t) i]; removeAt(i); return v; }
    T& first() { return this->front(); }
    const T& first() const { return this->front(); }
    T& last() { return this->back(); }
    const T& last() const { return this->back(); }
    bool contains(const T& v) const { return std::find(this->begin(), this->end(), v) != this->end(); }
    int indexOf(const T& v) const { auto it = std::find(this->begin(), this->end(), v); return it == this->end() ? -1 : (int) (it - this->begin()); }
    int removeAll(const T& v) { int n = 0; for (auto it = this->begin(); it != this->end();) { if (*it == v) { it = this->erase(it); ++n; } else ++it; } return n; }
    bool removeOne(const T& v) { auto it = std::find(this->begin(), this->end(), v); if (it == this->end()) return false; this->erase(it); return true; }
};
template <typename T> using QVector = QList<T>;
typedef QList<QString> QStringList;
template <typename T> class QListIterator {
public:
    QListIterator(const QList<T>& l) : l_(&l) {}
    bool hasNext() const { return i_ < l_->size(); }
    const T& next() { return (*l_)[i_++]; }
private:
    const QList<T>* l_; size_t i_ = 0;
};
template <typename T> class QMutableListIterator {
public:
    QMutableListIterator(QList<T>& l) : l_(&l) {}
    bool hasNext() const { return i_ < l_->size(); }
    T& next() { last_ = i_; return (*l_)[i_++]; }
    void remove() { l_->erase(l_->begin() + (long) last_); i_ = last_; }
    T& value() { return (*l_)[last_]; }
    void setValue(const T& v) { (*l_)[last_] = v; }
private:
    QList<T>* l_; size_t i_ = 0, last_ = 0;
};

template <typename K, typename V> class QHash {
public:
    QHash() {}
    QHash(std::initializer_list<std::pair<K, V>> l) { for (auto& p : l) m_[p.first] = p.second; }
    V& operator[](const K& k) { return m_[k]; }
    const V value(const K& k, const V& d = V()) const { auto it = m_.find(k); return it == m_.end() ? d : it->second; }
    bool contains(const K& k) const { return m_.find(k) != m_.end(); }
    int remove(const K& k) { return (int) m_.erase(k); }
    int count() const { return (int) m_.size(); }
    int size() const { return (int) m_.size(); }
    bool isEmpty() const { return m_.empty(); }
    void insert(const K& k, const V& v) { m_[k] = v; }
    void clear() { m_.clear(); }
    QList<K> keys() const { QList<K> r; for (auto& p : m_) r.push_back(p.first); return r; }
    QList<V> values() const { QList<V> r; for (auto& p : m_) r.push_back(p.second); return r; }
    typename std::map<K, V>::iterator begin() { return m_.begin(); }
    typename std::map<K, V>::iterator end() { return m_.end(); }
    typename std::map<K, V>::const_iterator begin() const { return m_.begin(); }
    typename std::map<K, V>::const_iterator end() const { return m_.end(); }
private:
    std::map<K, V> m_;
};
template <typename K, typename V> using QMap = QHash<K, V>;

class QVariant {
public:
    QVariant() {}
    QVariant(bool v) : d_(v ? 1.0 : 0.0), valid_(true) {}
    QVariant(int v) : d_(v), valid_(true) {}
    QVariant(float v) : d_(v), valid_(true) {}
    QVariant(double v) : d_(v), valid_(true) {}
    QVariant(qint64 v) : d_((double) v), valid_(true) {}
    QVariant(const QString& s) : s_(s), valid_(true) {}
    template <typename T> void setValue(const T& v) {
        if constexpr (std::is_pointer<T>::value) p_ = (void*) v; else d_ = (double) v;
        valid_ = true;
    }
    template <typename T> static QVariant fromValue(const T& v) { QVariant q; q.setValue<T>(v); return q; }
    template <typename T> T value() const {
        if constexpr (std::is_pointer<T>::value) return (T) p_; else return (T) d_;
    }
    void clear() { d_ = 0.0; s_ = QString(); p_ = nullptr; valid_ = false; }
    bool toBool() const { return d_ != 0.0; }
    int toInt() const { return (int) d_; }
    float toFloat() const { return (float) d_; }
    double toDouble() const { return d_; }
    QString toString() const { return s_; }
    bool isValid() const { return valid_; }
    bool operator==(const QVariant& o) const { return d_ == o.d_ && s_ == o.s_; }
private:
    double d_ = 0.0; QString s_; void* p_ = nullptr; bool valid_ = false;
};
typedef QList<QVariant> QVariantList;

class QThread;
class QObject {
public:
    explicit QObject(QObject* parent = nullptr) : parent_(parent) {}
    virtual ~QObject() {}
    QObject* parent() const { return parent_; }
    void setParent(QObject* p) { parent_ = p; }
    QThread* thread() const { return nullptr; }
    void moveToThread(QThread*) {}
    void deleteLater() {}
    void setObjectName(const QString&) {}
    template <typename... A> static bool connect(A&&...) { return true; }      // connections are never made: nothing on the pinned path is signal driven
    template <typename... A> static bool disconnect(A&&...) { return true; }
private:
    QObject* parent_;
};
template <typename T> inline T qobject_cast(QObject* o) { return dynamic_cast<T>(o); }
template <typename T> inline T qobject_cast(const QObject* o) { return dynamic_cast<T>(o); }

class QThread : public QObject {
public:
    explicit QThread(QObject* p = nullptr) : QObject(p) {}
    enum Priority { IdlePriority, LowestPriority, LowPriority, NormalPriority, HighPriority, HighestPriority, TimeCriticalPriority, InheritPriority };
    virtual void run() {}
    void start(Priority = InheritPriority) {}
    void quit() {}
    bool wait(unsigned long = ~0ul) { return true; }
    bool isRunning() const { return false; }
    static void msleep(unsigned long) {}
    static void usleep(unsigned long) {}
    static QThread* currentThread() { return nullptr; }
    void started() {}                                   // signals
    void finished() {}
};
struct QMetaObject {
    template <typename... A> static bool invokeMethod(A&&...) { return true; }       // queued invocations are never delivered: no event loop here
};
class QProcess : public QObject {
public:
    void start(const QString&, const QStringList& = QStringList()) {}
    bool waitForFinished(int = 30000) { return true; }
    QByteArray readAllStandardOutput() { return QByteArray(); }
};
class QCoreApplication : public QObject {
public:
    static QCoreApplication* instance() { static QCoreApplication app; return &app; }
    static void processEvents() {}
};
#define qApp (QCoreApplication::instance())

class QMutex {
public:
    void lock() { m_.lock(); }
    void unlock() { m_.unlock(); }
    bool tryLock(int = 0) { return m_.try_lock(); }
private:
    std::recursive_mutex m_;
};
class QMutexLocker {
public:
    explicit QMutexLocker(QMutex* m) : m_(m) { if (m_) m_->lock(); }
    ~QMutexLocker() { if (m_) m_->unlock(); }
private:
    QMutex* m_;
};
class QWaitCondition {
public:
    bool wait(QMutex*, unsigned long = ~0ul) { return true; }
    void wakeAll() {}
    void wakeOne() {}
};

class QTimer : public QObject {
public:
    explicit QTimer(QObject* p = nullptr) : QObject(p) {}
    void setInterval(int) {}
    void setSingleShot(bool) {}
    void setTimerType(Qt::TimerType) {}
    void start() {}
    void start(int) {}
    void stop() {}
    bool isActive() const { return false; }
    void timeout() {}                                   // signal
    template <typename... A> static void singleShot(A&&...) {}
};

// the test driver owns the clock (ClipAudioSourcePositionsModel drops positions that were not updated for a second)
extern "C" long long zl_ref_shim_now_msecs;
class QDateTime {
public:
    static qint64 currentMSecsSinceEpoch() { return zl_ref_shim_now_msecs; }
    static QDateTime currentDateTime() { return QDateTime(); }
    QString toString(Qt::DateFormat = Qt::TextDate) const { return QString("1970-01-01T00:00:00"); }
};
class QDir {
public:
    QDir() {}
    bool exists(const QString& path) const;          // oracle/ref/libzl_ref.cpp (std::filesystem)
    bool mkpath(const QString& path) const;
};

class QModelIndex {
public:
    QModelIndex() {}
    QModelIndex(int r, int c) : r_(r), c_(c) {}
    bool isValid() const { return r_ >= 0 && c_ >= 0; }
    int row() const { return r_; }
    int column() const { return c_; }
private:
    int r_ = -1, c_ = -1;
};
class QAbstractListModel : public QObject {
public:
    explicit QAbstractListModel(QObject* p = nullptr) : QObject(p) {}
    virtual QHash<int, QByteArray> roleNames() const { return QHash<int, QByteArray>(); }
    virtual int rowCount(const QModelIndex& parent) const = 0;
    virtual QVariant data(const QModelIndex& index, int role) const = 0;
    QModelIndex createIndex(int row, int col) const { return QModelIndex(row, col); }
    QModelIndex index(int row, int col = 0) const { return QModelIndex(row, col); }
    bool checkIndex(const QModelIndex& i) const { return i.isValid() && i.row() < rowCount(QModelIndex()); }
    void dataChanged(const QModelIndex&, const QModelIndex&, const QVector<int>& = QVector<int>()) {}   // signal
    void beginResetModel() {}
    void endResetModel() {}
};
