// registry_test.cpp -- the create / get / getDefault / erase / clear convention of the mirror classes
// (omega_b200/host/OmegaB200.h: Registry<T>), exercised on a class that needs no device, and a compile-time check that
// every mirror class carries it.  Host code only.
#include <cstdio>
#include <string>
#include <vector>

#include "OmegaB200.h"

using namespace OMEGA;

struct Named : Registry<Named> {
   std::string Name;
   int Value;
   Named(const std::string &N, int V) : Name(N), Value(V) { ++Alive; }
   ~Named() { --Alive; }
   static int Alive;
};
int Named::Alive = 0;
struct Plain : Registry<Plain> {
   int Value;
   explicit Plain(int V) : Value(V) {}
};

static_assert(std::is_base_of_v<Registry<HorzMesh>, HorzMesh> && std::is_base_of_v<Registry<Halo>, Halo> &&
                  std::is_base_of_v<Registry<OceanState>, OceanState> && std::is_base_of_v<Registry<AuxiliaryState>, AuxiliaryState> &&
                  std::is_base_of_v<Registry<Tendencies>, Tendencies> && std::is_base_of_v<Registry<TimeStepper>, TimeStepper>,
              "every mirror class follows the reference's ownership convention");

#define REQUIRE(c) \
   do { if (!(c)) { std::fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

int main() {
   REQUIRE(Named::getDefault() == nullptr && Named::get("Default") == nullptr);
   Named *D = Named::create("Default", 1);
   Named *B = Named::create("Second", 2);
   REQUIRE(D && B && D->Name == "Default" && B->Name == "Second" && Named::Alive == 2);
   REQUIRE(Named::getDefault() == D && Named::get("Second") == B && Named::get("Default") == D);
   REQUIRE(Named::create("Second", 3) == nullptr && Named::Alive == 2); // the name is taken
   Named::erase("Second");
   REQUIRE(Named::get("Second") == nullptr && Named::Alive == 1 && Named::getDefault() == D);
   Named::erase("NoSuchName");
   Named::erase("Default");
   REQUIRE(Named::getDefault() == nullptr && Named::Alive == 0);
   Named::create("Default", 4); Named::create("Other", 5);
   Named::clear();
   REQUIRE(Named::Alive == 0 && Named::getDefault() == nullptr && Named::get("Other") == nullptr);
   // a class whose constructor takes no name
   Plain *P = Plain::create("Default", 7);
   REQUIRE(P && P->Value == 7 && Plain::getDefault() == P);
   Plain::clear();
   // DeviceArray::slice (the subview of Tracers::getByIndex) over memory the test owns, and the tracer names
   {
      std::vector<double> Buf(2 * 3 * 4);
      for (size_t i = 0; i < Buf.size(); ++i) Buf[i] = (double)i;
      Array3DReal A("tracers", Buf.data(), {2, 3, 4});
      Array2DReal S1 = A.slice(1);
      REQUIRE(S1.data() == Buf.data() + 12 && S1.extent(0) == 3 && S1.extent(1) == 4 && S1.size() == 12 && S1.label() == "tracers");
      Array1DReal Row = S1.slice(2, "row");
      REQUIRE(Row.data() == Buf.data() + 12 + 8 && Row.extent(0) == 4 && Row.data()[3] == 23.0 && Row.label() == "row");
      TracerNames N;
      N.define({"Temperature", "Salinity", "Debug1"});
      I4 Idx = TracerNames::IndxInvalid;
      std::string Name;
      REQUIRE(N.getIndex(Idx, "Salinity") == 0 && Idx == 1 && N.getName(Name, 2) == 0 && Name == "Debug1" && N.size() == 3);
      REQUIRE(N.getIndex(Idx, "Oxygen") == -1 && Idx == 1 && N.getName(Name, 3) == -1 && N.getName(Name, -1) == -1);
   }
   std::printf("registry ok\n");
   return 0;
}
