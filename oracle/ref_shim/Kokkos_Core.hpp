// Kokkos_Core.hpp -- TEST INFRASTRUCTURE.  Minimal HOST stand-in for the part
// of the Kokkos API the reference's hot-path sources use, so
// that the reference's own, unmodified functor headers and sequencing sources
// compile here although the Kokkos submodule is missing from /root/reference
// (externals/ekat/extern/kokkos is empty).  Written from the public Kokkos
// API documentation; no Kokkos code is copied.  View = ref-counted
// LayoutRight array, parallel_for = (OpenMP) loop, host space only.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <iostream>
#include <map>
#include <vector>
#include <cstring>
#include <memory>
#include <string>
#include <type_traits>
#include <utility>

#define KOKKOS_FUNCTION
#define KOKKOS_INLINE_FUNCTION inline
#define KOKKOS_FORCEINLINE_FUNCTION inline
#define KOKKOS_LAMBDA [=]
#define KOKKOS_CLASS_LAMBDA [ =, *this ]

namespace Kokkos {

struct LayoutRight {};
struct LayoutLeft {};
struct ScratchSpace {};
struct OpenMP {
   using scratch_memory_space = ScratchSpace;
};
using Serial                = OpenMP;
using DefaultExecutionSpace = OpenMP;
struct HostSpace {
   using execution_space = OpenMP;
};
struct MemoryUnmanaged {};
inline void initialize(int &, char **) {}
inline void initialize() {}
inline void finalize() {}
inline void fence() {}
inline void fence(const std::string &) {}

// ---- View ------------------------------------------------------------------
template <class DT> struct ViewTraits {
   using value_type            = DT;
   static constexpr int rank   = 0;
};
template <class DT> struct ViewTraits<DT *> {
   using value_type          = typename ViewTraits<DT>::value_type;
   static constexpr int rank = ViewTraits<DT>::rank + 1;
};

// trailing compile-time extent: View<Real *[VecLength], ...> (TriDiagSolvers.h:37-39)
template <class DT, std::size_t N> struct ViewTraits<DT[N]> {
   using value_type                  = typename ViewTraits<DT>::value_type;
   static constexpr int rank         = ViewTraits<DT>::rank + 1;
   static constexpr std::size_t tail = N;
};
template <class T, class = void> struct StaticTail { static constexpr std::size_t value = 0; };
template <class T> struct StaticTail<T, std::void_t<decltype(ViewTraits<T>::tail)>> { static constexpr std::size_t value = ViewTraits<T>::tail; };

// per-team scratch memory: a bump allocator over a buffer the team loop owns (TeamPolicy::set_scratch_size)
struct ScratchArena {
   char *Base       = nullptr;
   std::size_t *Off = nullptr; // shared by the copies handed out by team_scratch()
   std::size_t Cap  = 0;
   void *alloc(std::size_t Bytes) const {
      const std::size_t a = (*Off + 15) & ~std::size_t(15);
      if (a + Bytes > Cap) { std::fprintf(stderr, "Kokkos stand-in: team scratch exhausted\n"); std::abort(); }
      *Off = a + Bytes;
      return Base + a;
   }
};

struct ALL_t {};
constexpr ALL_t ALL{};

template <class DataType, class... Props> class View {
 public:
   using data_type            = DataType;
   using value_type           = typename ViewTraits<DataType>::value_type;
   using non_const_value_type = std::remove_const_t<value_type>;
   // usable as a value and as a call, like Kokkos' integral constant (Halo.h:795 writes Array.rank())
   struct RankT {
      constexpr operator int() const { return ViewTraits<DataType>::rank; }
      constexpr int operator()() const { return ViewTraits<DataType>::rank; }
   };
   static constexpr RankT rank{};
   static constexpr int Rank         = ViewTraits<DataType>::rank;
   static constexpr bool is_hostspace = true;

   View() = default;
   template <class... Ns>
   explicit View(const std::string &Label, Ns... ns) : Label_(Label) {
      static_assert(sizeof...(Ns) == rank, "extent count must equal rank");
      const std::size_t e[] = {static_cast<std::size_t>(ns)..., 1};
      std::size_t total     = 1;
      for (int i = 0; i < rank; ++i) {
         Ext_[i] = e[i];
         total *= e[i];
      }
      // zero-initialised like a Kokkos::View
      Data_ = std::shared_ptr<non_const_value_type[]>(new non_const_value_type[total ? total : 1]());
      Ptr_  = Data_.get();
      Size_ = total;
   }
   // unmanaged view over caller memory
   template <class... Ns> View(value_type *Ptr, Ns... ns) : Label_("unmanaged"), Ptr_(Ptr) {
      const std::size_t e[] = {static_cast<std::size_t>(ns)..., 1};
      Size_                 = 1;
      for (int i = 0; i < rank; ++i) {
         Ext_[i] = e[i];
         Size_ *= e[i];
      }
   }

   // view in team scratch memory: dynamic extents as given, a trailing static extent from the type
   template <class... Ns> View(const ScratchArena &S, Ns... ns) : Label_("scratch") {
      const std::size_t e[] = {static_cast<std::size_t>(ns)..., 1};
      constexpr int nd      = static_cast<int>(sizeof...(Ns));
      Size_                 = 1;
      for (int i = 0; i < rank; ++i) {
         Ext_[i] = i < nd ? e[i] : StaticTail<DataType>::value;
         Size_ *= Ext_[i];
      }
      Ptr_ = static_cast<value_type *>(S.alloc(Size_ * sizeof(value_type)));
   }

   template <class... Is> value_type &operator()(Is... is) const {
      static_assert(sizeof...(Is) == rank, "index count must equal rank");
      const std::size_t idx[] = {static_cast<std::size_t>(is)..., 0};
      std::size_t off         = 0;
      for (int i = 0; i < rank; ++i) off = off * Ext_[i] + idx[i];
      return Ptr_[off];
   }
   template <class I> value_type &operator[](I i) const { return Ptr_[i]; }

   std::size_t extent(int i) const { return i < rank ? Ext_[i] : 1; }
   int extent_int(int i) const { return static_cast<int>(extent(i)); }
   std::size_t size() const { return Size_; }
   std::size_t span() const { return Size_; }
   value_type *data() const { return Ptr_; }
   const std::string &label() const { return Label_; }
   bool is_allocated() const { return Ptr_ != nullptr; }

 private:
   std::string Label_;
   std::shared_ptr<non_const_value_type[]> Data_;
   value_type *Ptr_ = nullptr;
   std::size_t Ext_[8] = {0, 0, 0, 0, 0, 0, 0, 0};
   std::size_t Size_   = 0;
};

template <class T> struct is_view : std::false_type {};
template <class D, class... P> struct is_view<View<D, P...>> : std::true_type {};
template <class D, class... P> struct is_view<const View<D, P...>> : std::true_type {};

// ---- make_pair / subview of a rank-1 view over a half-open index range (Halo.h:874-878, Halo.cpp:734-737) ----
using std::make_pair;
template <class D, class... P, class I1, class I2>
View<D, P...> subview(const View<D, P...> &V, const std::pair<I1, I2> &R) {
   static_assert(View<D, P...>::rank == 1, "stand-in: subview of rank-1 views only");
   return View<D, P...>(V.data() + R.first, static_cast<std::size_t>(R.second - R.first));
}

// ---- deep_copy / mirrors ------------------------------------------------------
template <class D, class... P, class S, std::enable_if_t<std::is_arithmetic_v<S>, int> = 0>
void deep_copy(const View<D, P...> &Dst, const S &Val) {
   std::fill(Dst.data(), Dst.data() + Dst.size(), static_cast<typename View<D, P...>::value_type>(Val));
}
template <class D1, class... P1, class D2, class... P2>
void deep_copy(const View<D1, P1...> &Dst, const View<D2, P2...> &Src) {
   if (Dst.data() != Src.data() && Dst.size())
      std::memcpy(Dst.data(), Src.data(), sizeof(typename View<D1, P1...>::value_type) * Dst.size());
}
template <class E, class Dst, class Src, std::enable_if_t<std::is_class_v<E> && !is_view<E>::value, int> = 0>
void deep_copy(const E &, const Dst &D, const Src &S) {
   deep_copy(D, S);
}
template <class Space, class D, class... P> View<D, P...> create_mirror_view_and_copy(const Space &, const View<D, P...> &Src) {
   return Src; // host == device
}
template <class D, class... P> View<D, P...> create_mirror_view(const View<D, P...> &Src) { return Src; }

// ---- policies ---------------------------------------------------------------
template <class T> struct IndexType {};
enum class Iterate { Default, Left, Right };
template <int N, Iterate Outer = Iterate::Default, Iterate Inner = Iterate::Default> struct Rank {
   static constexpr int rank = N;
};

template <class... Props> struct RangePolicy {
   long Begin, End;
   RangePolicy(long B, long E) : Begin(B), End(E) {}
};

template <class... Props> struct MDRangePolicy;
template <class Exec, int N, Iterate O, Iterate I, class... Rest> struct MDRangePolicy<Exec, Rank<N, O, I>, Rest...> {
   static constexpr int rank = N;
   long Lower[N], Upper[N];
   template <class A, class B> MDRangePolicy(const A (&L)[N], const B (&U)[N]) {
      for (int i = 0; i < N; ++i) {
         Lower[i] = L[i];
         Upper[i] = U[i];
      }
   }
};

// ---- hierarchical parallelism, serial: a league of one-thread teams (VertCoord.cpp:648-860) ----------------
struct PerTeamValue {
   std::size_t Bytes;
};
inline PerTeamValue PerTeam(std::size_t Bytes) { return {Bytes}; }
struct HostTeamMember {
   int League = 0;
   ScratchArena Scratch;
   const ScratchArena &team_scratch(int) const { return Scratch; }
   int league_rank() const { return League; }
   int team_rank() const { return 0; }
   int team_size() const { return 1; }
   void team_barrier() const {}
};
template <class Exec = OpenMP> struct TeamPolicy {
   using member_type = HostTeamMember;
   int LeagueSize    = 0;
   TeamPolicy() = default;
   std::size_t ScratchBytes = 0;
   TeamPolicy(int League, int /*TeamSize*/, int /*VectorLength*/ = 1) : LeagueSize(League) {}
   TeamPolicy &set_scratch_size(int, const PerTeamValue &V) {
      ScratchBytes = V.Bytes;
      return *this;
   }
};
struct TeamThreadRangeT {
   int Begin, End;
};
inline TeamThreadRangeT TeamThreadRange(const HostTeamMember &, int N) { return {0, N}; }
inline TeamThreadRangeT TeamThreadRange(const HostTeamMember &, int B, int E) { return {B, E}; }
template <class Exec, class F> void parallel_for(const std::string &, const TeamPolicy<Exec> &Pol, const F &Fun) {
   std::vector<char> buf(Pol.ScratchBytes + 64);
   for (int l = 0; l < Pol.LeagueSize; ++l) {
      std::size_t off = 0;
      HostTeamMember M;
      M.League  = l;
      M.Scratch = ScratchArena{buf.data(), &off, buf.size()};
      Fun(M);
   }
}
template <class Exec, class F> void parallel_for(const TeamPolicy<Exec> &Pol, const F &Fun) { parallel_for(std::string(), Pol, Fun); }
template <class F> void parallel_for(const TeamThreadRangeT &R, const F &Fun) {
   for (int i = R.Begin; i < R.End; ++i) Fun(i);
}
// the functor adds its contribution to the running value and, when IsFinal, uses it: one pass, in index order
template <class F, class C, class T> void team_scan_impl(const TeamThreadRangeT &R, const F &Fun, void (C::*)(int, T &, bool) const);
template <class F> void parallel_scan(const TeamThreadRangeT &R, const F &Fun) {
   // value type = second parameter of the functor's call operator
   using Traits = decltype(&F::operator());
   team_scan_impl(R, Fun, Traits{});
}
template <class F, class C, class T> void team_scan_impl(const TeamThreadRangeT &R, const F &Fun, void (C::*)(int, T &, bool) const) {
   T Acc{};
   for (int i = R.Begin; i < R.End; ++i) Fun(i, Acc, true);
}
template <class F, class T> void parallel_reduce(const TeamThreadRangeT &R, const F &Fun, T &Result) {
   T Acc{};
   for (int i = R.Begin; i < R.End; ++i) Fun(i, Acc);
   Result = Acc;
}
template <class F, class T1, class T2> void parallel_reduce(const TeamThreadRangeT &R, const F &Fun, T1 &R1, T2 &R2) {
   T1 A1{};
   T2 A2{};
   for (int i = R.Begin; i < R.End; ++i) Fun(i, A1, A2);
   R1 = A1;
   R2 = A2;
}

// ---- parallel_for ---------------------------------------------------------------
template <class... P, class F> void parallel_for(const std::string &, const RangePolicy<P...> &Pol, const F &Fun) {
#pragma omp parallel for schedule(static)
   for (long i = Pol.Begin; i < Pol.End; ++i) Fun(static_cast<int>(i));
}
template <class... P, class F> void parallel_for(const RangePolicy<P...> &Pol, const F &Fun) {
   parallel_for(std::string(), Pol, Fun);
}

template <class F, std::size_t... Is>
inline void call_md(const F &Fun, const std::array<int, sizeof...(Is)> &Idx, std::index_sequence<Is...>) {
   Fun(Idx[Is]...);
}

template <class... P, class F> void parallel_for(const std::string &, const MDRangePolicy<P...> &Pol, const F &Fun) {
   constexpr int N = MDRangePolicy<P...>::rank;
   long Inner      = 1;
   for (int d = 1; d < N; ++d) Inner *= (Pol.Upper[d] - Pol.Lower[d]);
   // parallel over the slowest index, row-major sweep of the rest
#pragma omp parallel for schedule(static)
   for (long i0 = Pol.Lower[0]; i0 < Pol.Upper[0]; ++i0) {
      std::array<int, N> Idx;
      Idx[0] = static_cast<int>(i0);
      for (long r = 0; r < Inner; ++r) {
         long rem = r;
         for (int d = N - 1; d >= 1; --d) {
            const long ext = Pol.Upper[d] - Pol.Lower[d];
            Idx[d]         = static_cast<int>(Pol.Lower[d] + rem % ext);
            rem /= ext;
         }
         call_md(Fun, Idx, std::make_index_sequence<N>{});
      }
   }
}
template <class... P, class F> void parallel_for(const MDRangePolicy<P...> &Pol, const F &Fun) {
   parallel_for(std::string(), Pol, Fun);
}

// ---- parallel_reduce (sum into a scalar, or a Max reducer; serial) ------------------
template <class T> struct Max {
   T &Ref;
   explicit Max(T &R) : Ref(R) {}
};
template <class T> struct Min {
   T &Ref;
   explicit Min(T &R) : Ref(R) {}
};
template <class... P, class F, class T, std::enable_if_t<std::is_arithmetic_v<T>, int> = 0>
void parallel_reduce(const std::string &, const RangePolicy<P...> &Pol, const F &Fun, T &Result) {
   T Acc = 0;
   for (long i = Pol.Begin; i < Pol.End; ++i) Fun(static_cast<int>(i), Acc);
   Result = Acc;
}
template <class... P, class F, class T> void parallel_reduce(const std::string &, const RangePolicy<P...> &Pol, const F &Fun, Max<T> R) {
   T Acc = R.Ref;
   for (long i = Pol.Begin; i < Pol.End; ++i) Fun(static_cast<int>(i), Acc);
   R.Ref = Acc;
}

// ---- math -------------------------------------------------------------------
using std::abs;
using std::ceil;
using std::log2;
using std::cos;
using std::fabs;
using std::sin;
using std::sqrt;
template <class T> inline T max(const T &a, const T &b) { return a < b ? b : a; }
template <class T> inline T min(const T &a, const T &b) { return b < a ? b : a; }

} // namespace Kokkos
