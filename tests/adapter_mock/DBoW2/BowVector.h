// Stand-in for thirdparty/DBoW2/DBoW2/BowVector.h (BowVector.h:56-66): a std::map<WordId, WordValue>.  TEST INFRASTRUCTURE.
#pragma once
#include <map>
namespace DBoW2 {
typedef unsigned int WordId;
typedef double WordValue;
class BowVector : public std::map<WordId, WordValue> {};
} // namespace DBoW2
