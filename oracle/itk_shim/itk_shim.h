// ITK stand-in used ONLY to compile the reference's hot-path headers for the
// oracle (test infrastructure, never linked into the product library).
// ITK is not installed in this image; the reference headers need a handful of
// ITK value types (glia_image.hxx:5-20).  This file supplies a minimal,
// independent implementation of exactly that surface: flat x-fastest buffers.
#ifndef GLIA_ORACLE_ITK_SHIM_H
#define GLIA_ORACLE_ITK_SHIM_H
#include <cstddef>
#include <cstdint>
#include <vector>
#include <memory>

namespace itk {

typedef long IndexValueType;
typedef unsigned long SizeValueType;
typedef long OffsetValueType;

template <unsigned int D>
struct Index {
  static const unsigned int Dimension = D;
  IndexValueType m_Index[D];
  IndexValueType& operator[](unsigned int i) { return m_Index[i]; }
  IndexValueType operator[](unsigned int i) const { return m_Index[i]; }
  void Fill(IndexValueType v) { for (unsigned int i = 0; i < D; ++i) m_Index[i] = v; }
  const IndexValueType* begin() const { return m_Index; }
  const IndexValueType* end() const { return m_Index + D; }
  bool operator==(Index const& o) const {
    for (unsigned int i = 0; i < D; ++i) if (m_Index[i] != o.m_Index[i]) return false;
    return true;
  }
  bool operator!=(Index const& o) const { return !(*this == o); }
};

template <unsigned int D>
struct Size {
  SizeValueType m_Size[D];
  SizeValueType& operator[](unsigned int i) { return m_Size[i]; }
  SizeValueType operator[](unsigned int i) const { return m_Size[i]; }
  void Fill(SizeValueType v) { for (unsigned int i = 0; i < D; ++i) m_Size[i] = v; }
};

template <typename T, unsigned int D>
struct Point {
  T m_Data[D];
  Point() { for (unsigned int i = 0; i < D; ++i) m_Data[i] = T(); }
  T& operator[](unsigned int i) { return m_Data[i]; }
  T const& operator[](unsigned int i) const { return m_Data[i]; }
  T* Begin() { return m_Data; }
  T const* Begin() const { return m_Data; }
  void Fill(T const& v) { for (unsigned int i = 0; i < D; ++i) m_Data[i] = v; }
};

template <unsigned int D>
class ImageRegion {
 public:
  static const unsigned int ImageDimension = D;
  typedef Index<D> IndexType;
  typedef Size<D> SizeType;
  ImageRegion() { m_Index.Fill(0); m_Size.Fill(0); }
  virtual ~ImageRegion() {}
  IndexType const& GetIndex() const { return m_Index; }
  SizeType const& GetSize() const { return m_Size; }
  IndexType& GetModifiableIndex() { return m_Index; }
  SizeType& GetModifiableSize() { return m_Size; }
  void SetIndex(IndexType const& i) { m_Index = i; }
  void SetSize(SizeType const& s) { m_Size = s; }
  bool IsInside(IndexType const& p) const {
    for (unsigned int i = 0; i < D; ++i) {
      if (p[i] < m_Index[i]) return false;
      if (p[i] >= m_Index[i] + (IndexValueType)m_Size[i]) return false;
    }
    return true;
  }
 private:
  IndexType m_Index;
  SizeType m_Size;
};

template <typename T>
class SmartPointer {
 public:
  typedef T ObjectType;
  SmartPointer() {}
  SmartPointer(std::nullptr_t) {}
  explicit SmartPointer(std::shared_ptr<T> const& p) : m_P(p) {}
  bool IsNull() const { return !m_P; }
  bool IsNotNull() const { return (bool)m_P; }
  T* operator->() const { return m_P.get(); }
  T* GetPointer() const { return m_P.get(); }
  T& operator*() const { return *m_P; }
 private:
  std::shared_ptr<T> m_P;
};

template <typename T, unsigned int D>
class Image {
 public:
  static const unsigned int ImageDimension = D;
  typedef T PixelType;
  typedef SmartPointer<Image> Pointer;
  typedef ImageRegion<D> RegionType;
  typedef Index<D> IndexType;
  static Pointer New() { return Pointer(std::make_shared<Image>()); }
  void SetRegions(RegionType const& r) { m_Region = r; }
  void Allocate() {
    std::size_t n = 1;
    for (unsigned int i = 0; i < D; ++i) { m_Stride[i] = n; n *= m_Region.GetSize()[i]; }
    m_Buf.assign(n, T());
  }
  void FillBuffer(T const& v) { m_Buf.assign(m_Buf.size(), v); }
  RegionType const& GetRequestedRegion() const { return m_Region; }
  RegionType const& GetLargestPossibleRegion() const { return m_Region; }
  std::size_t Offset(IndexType const& p) const {
    std::size_t o = 0;
    for (unsigned int i = 0; i < D; ++i)
      o += (std::size_t)(p[i] - m_Region.GetIndex()[i]) * m_Stride[i];
    return o;
  }
  T const& GetPixel(IndexType const& p) const { return m_Buf[Offset(p)]; }
  void SetPixel(IndexType const& p, T const& v) { m_Buf[Offset(p)] = v; }
  T* GetBufferPointer() { return m_Buf.data(); }
  T const* GetBufferPointer() const { return m_Buf.data(); }
  std::size_t GetBufferSize() const { return m_Buf.size(); }
 private:
  RegionType m_Region;
  std::size_t m_Stride[D];
  std::vector<T> m_Buf;
};

// Walks the buffer in storage order (dimension 0 fastest), tracking the index.
template <typename TImage>
class ImageRegionConstIteratorWithIndex {
 public:
  typedef typename TImage::IndexType IndexType;
  typedef typename TImage::PixelType PixelType;
  template <typename TPtr, typename TRegion>
  ImageRegionConstIteratorWithIndex(TPtr const& image, TRegion const& region)
      : m_Image(&*image), m_Region(region), m_Off(0) {
    m_Index = m_Region.GetIndex();
    m_N = 1;
    for (unsigned int i = 0; i < TImage::ImageDimension; ++i) m_N *= m_Region.GetSize()[i];
  }
  bool IsAtEnd() const { return m_Off >= m_N; }
  ImageRegionConstIteratorWithIndex& operator++() {
    ++m_Off;
    for (unsigned int i = 0; i < TImage::ImageDimension; ++i) {
      if (++m_Index[i] < m_Region.GetIndex()[i] + (IndexValueType)m_Region.GetSize()[i]) break;
      if (i + 1 < TImage::ImageDimension) m_Index[i] = m_Region.GetIndex()[i];
    }
    return *this;
  }
  PixelType const& Value() const { return m_Image->GetPixel(m_Index); }
  PixelType Get() const { return m_Image->GetPixel(m_Index); }
  IndexType const& GetIndex() const { return m_Index; }
 private:
  TImage const* m_Image;
  typename TImage::RegionType m_Region;
  IndexType m_Index;
  std::size_t m_Off, m_N;
};

// Names that glia_image.hxx aliases but the hot path never instantiates.
template <typename T, unsigned int D> class VectorImage {};
template <typename T> class RGBPixel {};
template <typename TImage> class ImageRegionIterator {};
template <typename TImage> class ImageRegionConstIterator {};
template <typename TImage> class ImageRegionIteratorWithIndex {};
template <typename TImage> class ImageLinearIteratorWithIndex {};
template <typename TImage> class ImageLinearConstIteratorWithIndex {};
template <typename TImage> class ImageSliceIteratorWithIndex {};
template <typename TImage> class ImageSliceConstIteratorWithIndex {};

}  // namespace itk
#endif
