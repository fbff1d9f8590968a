// Minimal stand-ins for the Qt value classes the SME headers mention, with Qt's
// own signatures — enough for a -fsyntax-only pass, nothing is ever linked or run.
#pragma once
#include <cstddef>
#include <cstdint>
#include <initializer_list>
#include <string>
#include <vector>
using QRgb = unsigned int;
using qsizetype = std::ptrdiff_t;
using qreal = double;
inline int qRed(QRgb) { return 0; }
inline int qGreen(QRgb) { return 0; }
inline int qBlue(QRgb) { return 0; }
inline int qAlpha(QRgb) { return 0; }
inline QRgb qRgb(int, int, int) { return 0; }
inline QRgb qRgba(int, int, int, int) { return 0; }
class QByteArray {
public:
  const char *constData() const;
  const char *data() const;
  qsizetype size() const;
};
class QChar {
public:
  QChar() = default;
  QChar(char);
};
class QString {
public:
  QString() = default;
  QString(const char *);
  QString(const QChar &);
  static QString fromStdString(const std::string &);
  static QString number(double, char = 'g', int = 6);
  static QString number(int, int = 10);
  static QString number(long, int = 10);
  static QString number(unsigned long, int = 10);
  std::string toStdString() const;
  QByteArray toUtf8() const;
  QByteArray toLatin1() const;
  bool isEmpty() const;
  qsizetype size() const;
  qsizetype length() const;
  QString arg(const QString &) const;
  QString arg(double) const;
  QString arg(int) const;
  QString &append(const QString &);
  QString &operator+=(const QString &);
  double toDouble(bool * = nullptr) const;
  int toInt(bool * = nullptr) const;
  bool contains(const QString &) const;
  bool startsWith(const QString &) const;
  bool endsWith(const QString &) const;
  QString left(qsizetype) const;
  QString right(qsizetype) const;
  QString mid(qsizetype, qsizetype = -1) const;
  void clear();
  friend bool operator==(const QString &, const QString &);
  friend bool operator!=(const QString &, const QString &);
  friend bool operator<(const QString &, const QString &);
  friend QString operator+(const QString &, const QString &);
};
template <typename T> class QList {
public:
  using value_type = T;
  using iterator = T *;
  using const_iterator = const T *;
  using size_type = qsizetype;
  using reference = T &;
  using const_reference = const T &;
  QList() = default;
  QList(std::initializer_list<T>);
  explicit QList(qsizetype);
  QList(qsizetype, const T &);
  qsizetype size() const;
  qsizetype count() const;
  bool isEmpty() const;
  bool empty() const;
  T &operator[](qsizetype);
  const T &operator[](qsizetype) const;
  const T &at(qsizetype) const;
  void push_back(const T &);
  void append(const T &);
  void reserve(qsizetype);
  void resize(qsizetype);
  void clear();
  T &front();
  const T &front() const;
  T &back();
  const T &back() const;
  iterator begin();
  iterator end();
  const_iterator begin() const;
  const_iterator end() const;
  const_iterator cbegin() const;
  const_iterator cend() const;
  qsizetype indexOf(const T &) const;
  bool contains(const T &) const;
  T *data();
  const T *data() const;
  friend bool operator==(const QList &, const QList &) { return true; }
};
template <typename T> using QVector = QList<T>;
class QStringList : public QList<QString> {
public:
  using QList<QString>::QList;
  QString join(const QString &) const;
};
class QPoint {
public:
  constexpr QPoint() = default;
  constexpr QPoint(int x, int y) : xp(x), yp(y) {}
  constexpr int x() const { return xp; }
  constexpr int y() const { return yp; }
  int &rx() { return xp; }
  int &ry() { return yp; }
  void setX(int);
  void setY(int);
  friend constexpr bool operator==(const QPoint &a, const QPoint &b) {
    return a.xp == b.xp && a.yp == b.yp;
  }
  friend constexpr bool operator!=(const QPoint &a, const QPoint &b) {
    return !(a == b);
  }
  friend constexpr QPoint operator+(const QPoint &a, const QPoint &b) {
    return {a.xp + b.xp, a.yp + b.yp};
  }
  friend constexpr QPoint operator-(const QPoint &a, const QPoint &b) {
    return {a.xp - b.xp, a.yp - b.yp};
  }
  QPoint &operator+=(const QPoint &);
  QPoint &operator-=(const QPoint &);
  friend constexpr QPoint operator*(int c, const QPoint &a) {
    return {a.xp * c, a.yp * c};
  }
  friend constexpr QPoint operator*(const QPoint &a, int c) {
    return {a.xp * c, a.yp * c};
  }

private:
  int xp{0}, yp{0};
};
class QPointF {
public:
  constexpr QPointF() = default;
  constexpr QPointF(double x, double y) : xp(x), yp(y) {}
  constexpr QPointF(const QPoint &p) : xp(p.x()), yp(p.y()) {}
  constexpr double x() const { return xp; }
  constexpr double y() const { return yp; }
  double &rx() { return xp; }
  double &ry() { return yp; }
  void setX(double);
  void setY(double);
  QPoint toPoint() const;
  friend constexpr bool operator==(const QPointF &a, const QPointF &b) {
    return a.xp == b.xp && a.yp == b.yp;
  }
  friend constexpr QPointF operator+(const QPointF &a, const QPointF &b) {
    return {a.xp + b.xp, a.yp + b.yp};
  }
  friend constexpr QPointF operator-(const QPointF &a, const QPointF &b) {
    return {a.xp - b.xp, a.yp - b.yp};
  }
  friend constexpr QPointF operator*(const QPointF &a, double c) {
    return {a.xp * c, a.yp * c};
  }
  friend constexpr QPointF operator*(double c, const QPointF &a) {
    return {a.xp * c, a.yp * c};
  }
  friend constexpr QPointF operator/(const QPointF &a, double c) {
    return {a.xp / c, a.yp / c};
  }
  QPointF &operator+=(const QPointF &);
  QPointF &operator-=(const QPointF &);
  QPointF &operator*=(double);
  QPointF &operator/=(double);

private:
  double xp{0}, yp{0};
};
class QSize {
public:
  constexpr QSize() = default;
  constexpr QSize(int w, int h) : wd(w), ht(h) {}
  constexpr int width() const { return wd; }
  constexpr int height() const { return ht; }
  int &rwidth() { return wd; }
  int &rheight() { return ht; }
  void setWidth(int);
  void setHeight(int);
  bool isEmpty() const;
  friend constexpr bool operator==(const QSize &a, const QSize &b) {
    return a.wd == b.wd && a.ht == b.ht;
  }
  friend constexpr bool operator!=(const QSize &a, const QSize &b) {
    return !(a == b);
  }
  friend constexpr QSize operator*(const QSize &a, int c) {
    return {a.wd * c, a.ht * c};
  }

private:
  int wd{-1}, ht{-1};
};
class QSizeF {
public:
  constexpr QSizeF() = default;
  constexpr QSizeF(double w, double h) : wd(w), ht(h) {}
  constexpr QSizeF(const QSize &s) : wd(s.width()), ht(s.height()) {}
  constexpr double width() const { return wd; }
  constexpr double height() const { return ht; }
  double &rwidth() { return wd; }
  double &rheight() { return ht; }

private:
  double wd{-1}, ht{-1};
};
class QColor {
public:
  QColor() = default;
  QColor(int, int, int, int = 255);
  QColor(QRgb);
  QRgb rgb() const;
  QRgb rgba() const;
  int red() const;
  int green() const;
  int blue() const;
};
namespace Qt {
enum AspectRatioMode { IgnoreAspectRatio, KeepAspectRatio };
enum TransformationMode { FastTransformation, SmoothTransformation };
enum GlobalColor { black, white, transparent };
enum ImageConversionFlag {
  AutoColor = 0,
  ThresholdDither = 0x10,
  ThresholdAlphaDither = 0x04,
  AvoidDither = 0x80,
  NoOpaqueDetection = 0x100
};
constexpr ImageConversionFlag operator|(ImageConversionFlag a, ImageConversionFlag b) {
  return static_cast<ImageConversionFlag>(static_cast<int>(a) | static_cast<int>(b));
}
} // namespace Qt
class QImage {
public:
  enum Format {
    Format_Invalid,
    Format_Mono,
    Format_Indexed8,
    Format_RGB32,
    Format_ARGB32,
    Format_ARGB32_Premultiplied
  };
  QImage() = default;
  QImage(int, int, Format);
  QImage(const QSize &, Format);
  explicit QImage(const QString &);
  int width() const;
  int height() const;
  QSize size() const;
  bool isNull() const;
  Format format() const;
  QRgb pixel(int, int) const;
  QRgb pixel(const QPoint &) const;
  int pixelIndex(int, int) const;
  int pixelIndex(const QPoint &) const;
  void setPixel(int, int, unsigned int);
  void setPixel(const QPoint &, unsigned int);
  void fill(unsigned int);
  void fill(Qt::GlobalColor);
  bool valid(int, int) const;
  bool valid(const QPoint &) const;
  QImage convertToFormat(Format, Qt::ImageConversionFlag = Qt::AutoColor) const;
  QImage convertToFormat(Format, const QList<QRgb> &, Qt::ImageConversionFlag = Qt::AutoColor) const;
  QImage copy() const;
  QImage scaled(int, int, Qt::AspectRatioMode = Qt::IgnoreAspectRatio,
                Qt::TransformationMode = Qt::FastTransformation) const;
  QImage scaled(const QSize &, Qt::AspectRatioMode = Qt::IgnoreAspectRatio,
                Qt::TransformationMode = Qt::FastTransformation) const;
  QImage mirrored(bool = false, bool = true) const;
  QList<QRgb> colorTable() const;
  void setColorTable(const QList<QRgb> &);
  int colorCount() const;
  void setColor(int, QRgb);
  QRgb color(int) const;
  bool load(const QString &);
  bool save(const QString &) const;
  unsigned char *bits();
  const unsigned char *constBits() const;
  unsigned char *scanLine(int);
  const unsigned char *constScanLine(int) const;
  friend bool operator==(const QImage &, const QImage &);
};
