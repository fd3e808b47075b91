// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h header; parity unpinned).
//
// Bin-boundary search and value->bin mapping: restates upstream BinMapper::FindBin / FindBinWithZeroAsOneBin /
// GreedyFindBin (src/io/bin.cpp) and BinMapper::ValueToBin (include/LightGBM/bin.h), SURVEY.md A.1-A.4.
// Reference trigger: LGBM_DatasetCreateFromSampledColumn at
// lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/dataset/ReferenceDatasetUtils.scala:38-47
// with the column-major non-zero sample built by dataset/SampledData.scala:26-96.
#pragma once
#include <algorithm>
#include <unordered_map>

#include "orc_common.h"

namespace orc {

inline std::vector<double> GreedyFindBin(const double* distinct, const int* counts, int num_distinct, int max_bin,
                                         size_t total_cnt, int min_data_in_bin) {
  std::vector<double> bounds;
  if (num_distinct <= max_bin) {
    int cur = 0;
    for (int i = 0; i < num_distinct - 1; ++i) {
      cur += counts[i];
      if (cur >= min_data_in_bin) {
        double val = UpperBoundOf((distinct[i] + distinct[i + 1]) / 2.0);
        if (bounds.empty() || !EqualOrdered(bounds.back(), val)) {
          bounds.push_back(val);
          cur = 0;
        }
      }
    }
    bounds.push_back(std::numeric_limits<double>::infinity());
    return bounds;
  }
  if (min_data_in_bin > 0) {
    max_bin = std::min(max_bin, static_cast<int>(total_cnt / min_data_in_bin));
    max_bin = std::max(max_bin, 1);
  }
  double mean_bin_size = static_cast<double>(total_cnt) / max_bin;
  int rest_bin_cnt = max_bin;
  int rest_sample_cnt = static_cast<int>(total_cnt);
  std::vector<char> is_big(num_distinct, 0);
  for (int i = 0; i < num_distinct; ++i) {
    if (counts[i] >= mean_bin_size) {
      is_big[i] = 1;
      --rest_bin_cnt;
      rest_sample_cnt -= counts[i];
    }
  }
  mean_bin_size = static_cast<double>(rest_sample_cnt) / rest_bin_cnt;
  std::vector<double> upper(max_bin, std::numeric_limits<double>::infinity());
  std::vector<double> lower(max_bin, std::numeric_limits<double>::infinity());
  int bin_cnt = 0;
  lower[bin_cnt] = distinct[0];
  int cur = 0;
  for (int i = 0; i < num_distinct - 1; ++i) {
    if (!is_big[i]) rest_sample_cnt -= counts[i];
    cur += counts[i];
    if (is_big[i] || cur >= mean_bin_size ||
        (is_big[i + 1] && cur >= std::max(1.0, mean_bin_size * 0.5f))) {
      upper[bin_cnt] = distinct[i];
      ++bin_cnt;
      lower[bin_cnt] = distinct[i + 1];
      if (bin_cnt >= max_bin - 1) break;
      cur = 0;
      if (!is_big[i]) {
        --rest_bin_cnt;
        mean_bin_size = rest_sample_cnt / static_cast<double>(rest_bin_cnt);
      }
    }
  }
  ++bin_cnt;
  for (int i = 0; i < bin_cnt - 1; ++i) {
    double val = UpperBoundOf((upper[i] + lower[i + 1]) / 2.0);
    if (bounds.empty() || !EqualOrdered(bounds.back(), val)) bounds.push_back(val);
  }
  bounds.push_back(std::numeric_limits<double>::infinity());
  return bounds;
}

inline std::vector<double> FindBinWithZeroAsOneBin(const double* distinct, const int* counts, int num_distinct,
                                                   int max_bin, size_t total_sample_cnt, int min_data_in_bin) {
  std::vector<double> bounds;
  int left_cnt_data = 0, cnt_zero = 0, right_cnt_data = 0;
  for (int i = 0; i < num_distinct; ++i) {
    if (distinct[i] <= -kZeroThreshold) left_cnt_data += counts[i];
    else if (distinct[i] > kZeroThreshold) right_cnt_data += counts[i];
    else cnt_zero += counts[i];
  }
  int left_cnt = -1;
  for (int i = 0; i < num_distinct; ++i) {
    if (distinct[i] > -kZeroThreshold) { left_cnt = i; break; }
  }
  if (left_cnt < 0) left_cnt = num_distinct;
  if (left_cnt > 0 && max_bin > 1) {
    int left_max_bin = static_cast<int>(static_cast<double>(left_cnt_data) / (total_sample_cnt - cnt_zero) * (max_bin - 1));
    left_max_bin = std::max(1, left_max_bin);
    bounds = GreedyFindBin(distinct, counts, left_cnt, left_max_bin, left_cnt_data, min_data_in_bin);
    if (!bounds.empty()) bounds.back() = -kZeroThreshold;
  }
  int right_start = -1;
  for (int i = left_cnt; i < num_distinct; ++i) {
    if (distinct[i] > kZeroThreshold) { right_start = i; break; }
  }
  int right_max_bin = max_bin - 1 - static_cast<int>(bounds.size());
  if (right_start >= 0 && right_max_bin > 0) {
    auto right = GreedyFindBin(distinct + right_start, counts + right_start, num_distinct - right_start,
                               right_max_bin, right_cnt_data, min_data_in_bin);
    bounds.push_back(kZeroThreshold);
    bounds.insert(bounds.end(), right.begin(), right.end());
  } else {
    bounds.push_back(std::numeric_limits<double>::infinity());
  }
  return bounds;
}

struct BinMapper {
  int num_bin = 1;
  int missing_type = MISSING_NONE;
  bool is_trivial = true;
  double sparse_rate = 1.0;
  std::vector<double> bin_upper_bound;
  double min_val = 0.0, max_val = 0.0;
  uint32_t default_bin = 0;
  uint32_t most_freq_bin = 0;
  // categorical features (BinType::CategoricalBin): bin 0 collects NaN, negative, rare and unseen categories; the kept categories
  // follow by falling sample count
  bool categorical = false;
  std::vector<int> bin_2_categorical;
  std::unordered_map<int, unsigned int> categorical_2_bin;

  uint32_t ValueToBin(double value) const {
    if (categorical) {
      if (std::isnan(value)) return 0;
      if (!(value < 2147483648.0)) return 0;      // beyond int: the cast upstream makes is undefined there; treated as unseen
      const int iv = static_cast<int>(value);
      if (iv < 0) return 0;
      auto it = categorical_2_bin.find(iv);
      return it == categorical_2_bin.end() ? 0 : it->second;
    }
    if (std::isnan(value)) {
      if (missing_type == MISSING_NAN) return num_bin - 1;
      value = 0.0;
    }
    int l = 0, r = num_bin - 1;
    if (missing_type == MISSING_NAN) r -= 1;
    while (l < r) {
      int m = (r + l - 1) / 2;
      if (value <= bin_upper_bound[m]) r = m; else l = m + 1;
    }
    return l;
  }

  // values: the sampled NON-ZERO values of the feature (copied by the caller; sorted in place here).
  void FindBin(double* values, int num_sample_values, size_t total_sample_cnt, int max_bin, int min_data_in_bin,
               int min_split_data, bool pre_filter, bool use_missing, bool zero_as_missing, bool is_categorical = false) {
    categorical = is_categorical;
    int na_cnt = 0;
    int tmp = 0;
    for (int i = 0; i < num_sample_values; ++i) {
      if (!std::isnan(values[i])) values[tmp++] = values[i];
    }
    if (!use_missing) missing_type = MISSING_NONE;
    else if (zero_as_missing) missing_type = MISSING_ZERO;
    else if (tmp == num_sample_values) missing_type = MISSING_NONE;
    else { missing_type = MISSING_NAN; na_cnt = num_sample_values - tmp; }
    num_sample_values = tmp;
    default_bin = 0;
    int zero_cnt = static_cast<int>(total_sample_cnt - num_sample_values - na_cnt);
    std::vector<double> distinct;
    std::vector<int> counts;
    std::stable_sort(values, values + num_sample_values);
    if (num_sample_values == 0 || (values[0] > 0.0f && zero_cnt > 0)) {
      distinct.push_back(0.0);
      counts.push_back(zero_cnt);
    }
    if (num_sample_values > 0) {
      distinct.push_back(values[0]);
      counts.push_back(1);
    }
    for (int i = 1; i < num_sample_values; ++i) {
      if (!EqualOrdered(values[i - 1], values[i])) {
        if (values[i - 1] < 0.0f && values[i] > 0.0f) {
          distinct.push_back(0.0);
          counts.push_back(zero_cnt);
        }
        distinct.push_back(values[i]);
        counts.push_back(1);
      } else {
        distinct.back() = values[i];   // keep the larger representative
        ++counts.back();
      }
    }
    if (num_sample_values > 0 && values[num_sample_values - 1] < 0.0f && zero_cnt > 0) {
      distinct.push_back(0.0);
      counts.push_back(zero_cnt);
    }
    min_val = distinct.front();
    max_val = distinct.back();
    int nd = static_cast<int>(distinct.size());
    std::vector<int> cnt_in_bin;
    if (categorical) {
      // BinMapper::FindBin, bin_type == CategoricalBin: values become ints (negative ones count as missing), the categories are
      // taken by falling count until 99 % of the sample is covered AND max_bin bins exist; categories seen fewer than
      // min_data_in_bin times end the list (except the first two)
      std::vector<int> dvi, cti;
      for (int i = 0; i < nd; ++i) {
        const int val = distinct[i] < 2147483648.0 ? static_cast<int>(distinct[i]) : -1;      // beyond int: counted as missing
        if (val < 0) na_cnt += counts[i];
        else if (dvi.empty() || val != dvi.back()) { dvi.push_back(val); cti.push_back(counts[i]); }
        else cti.back() += counts[i];
      }
      num_bin = 0;
      const int rest_cnt = static_cast<int>(total_sample_cnt - na_cnt);
      if (rest_cnt > 0) {
        std::vector<int> order(dvi.size());
        for (size_t i = 0; i < order.size(); ++i) order[i] = static_cast<int>(i);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cti[a] > cti[b]; });     // SortForPair(..., is_reverse)
        const int cut_cnt = RoundInt((total_sample_cnt - na_cnt) * 0.99f);
        size_t cur_cat = 0;
        int used_cnt = 0;
        int distinct_cnt = static_cast<int>(dvi.size());
        if (na_cnt > 0) ++distinct_cnt;
        max_bin = std::min(distinct_cnt, max_bin);
        bin_2_categorical.assign(1, -1);
        categorical_2_bin.clear();
        categorical_2_bin[-1] = 0;
        cnt_in_bin.assign(1, 0);
        num_bin = 1;
        while (cur_cat < order.size() && (used_cnt < cut_cnt || num_bin < max_bin)) {
          const int o = order[cur_cat];
          if (cti[o] < min_data_in_bin && cur_cat > 1) break;
          bin_2_categorical.push_back(dvi[o]);
          categorical_2_bin[dvi[o]] = static_cast<unsigned int>(num_bin);
          used_cnt += cti[o];
          cnt_in_bin.push_back(cti[o]);
          ++num_bin;
          ++cur_cat;
        }
        // MissingType::None here says that every category seen has a bin of its own
        missing_type = (cur_cat == order.size() && na_cnt == 0) ? MISSING_NONE : MISSING_NAN;
        cnt_in_bin[0] = static_cast<int>(total_sample_cnt - used_cnt);
      }
      bin_upper_bound.clear();
    } else if (missing_type == MISSING_ZERO) {
      bin_upper_bound = FindBinWithZeroAsOneBin(distinct.data(), counts.data(), nd, max_bin, total_sample_cnt, min_data_in_bin);
      if (bin_upper_bound.size() == 2) missing_type = MISSING_NONE;
    } else if (missing_type == MISSING_NONE) {
      bin_upper_bound = FindBinWithZeroAsOneBin(distinct.data(), counts.data(), nd, max_bin, total_sample_cnt, min_data_in_bin);
    } else {
      bin_upper_bound = FindBinWithZeroAsOneBin(distinct.data(), counts.data(), nd, max_bin - 1, total_sample_cnt - na_cnt, min_data_in_bin);
      bin_upper_bound.push_back(std::numeric_limits<double>::quiet_NaN());
    }
    if (!categorical) {
      num_bin = static_cast<int>(bin_upper_bound.size());
      cnt_in_bin.assign(num_bin, 0);
      int i_bin = 0;
      for (int i = 0; i < nd; ++i) {
        while (distinct[i] > bin_upper_bound[i_bin] && i_bin < num_bin - 1) ++i_bin;
        cnt_in_bin[i_bin] += counts[i];
      }
      if (missing_type == MISSING_NAN) cnt_in_bin[num_bin - 1] = na_cnt;
    }
    is_trivial = num_bin <= 1;
    if (!is_trivial && pre_filter) {
      // NeedFilter: no cut leaves >= min_split_data samples on both sides (categorical: no single bin against the rest, looked at
      // only when there are at most two bins)
      bool need = true;
      int sum_left = 0;
      int total = static_cast<int>(total_sample_cnt);
      if (!categorical) {
        for (size_t i = 0; i + 1 < cnt_in_bin.size(); ++i) {
          sum_left += cnt_in_bin[i];
          if (sum_left >= min_split_data && total - sum_left >= min_split_data) { need = false; break; }
        }
      } else if (cnt_in_bin.size() <= 2) {
        for (size_t i = 0; i + 1 < cnt_in_bin.size(); ++i) {
          sum_left = cnt_in_bin[i];
          if (sum_left >= min_split_data && total - sum_left >= min_split_data) { need = false; break; }
        }
      } else {
        need = false;
      }
      if (need) is_trivial = true;
    }
    if (!is_trivial) {
      default_bin = ValueToBin(0);
      int arg = 0;
      for (int i = 1; i < num_bin; ++i) if (cnt_in_bin[i] > cnt_in_bin[arg]) arg = i;
      most_freq_bin = static_cast<uint32_t>(arg);
      double max_sparse_rate = static_cast<double>(cnt_in_bin[most_freq_bin]) / total_sample_cnt;
      if (most_freq_bin != default_bin && max_sparse_rate < kSparseThreshold) most_freq_bin = default_bin;
      sparse_rate = static_cast<double>(cnt_in_bin[most_freq_bin]) / total_sample_cnt;
    } else {
      sparse_rate = 1.0;
    }
  }
};

}  // namespace orc
