// filter_stage_test.cpp -- host logic of ColumnData::FilterAggregate without a GPU: how a pushed TableFilter becomes the
// chain of adacom_scan_select calls (ColumnData::BuildFilterStages).  Mirrors the recursion of
// ColumnSegment::FilterSelection (src/storage/table/column_segment.cpp:685-844): leaves and same-kind nesting flatten
// into one call, OR groups under an AND refine the selection one call each, an AND group under an OR is refused.
#include "../adacom_host.hpp"

#include <cstdio>
#include <cstdlib>

using namespace adacom_host;

#define CHECK(cond)                                                       \
	do {                                                                  \
		if (!(cond)) {                                                    \
			printf("CHECK failed at line %d: %s\n", __LINE__, #cond);     \
			exit(1);                                                      \
		}                                                                 \
	} while (0)

static unique_ptr<TableFilter> Cmp(ExpressionType t, int64_t c) {
	return unique_ptr<TableFilter>(new ConstantFilter(t, c));
}

int main() {
	// a leaf: one AND stage with one child; the six comparison types map to adacom_cmp
	const ExpressionType types[6] = {ExpressionType::COMPARE_EQUAL,           ExpressionType::COMPARE_NOTEQUAL,
	                                 ExpressionType::COMPARE_LESSTHAN,        ExpressionType::COMPARE_GREATERTHAN,
	                                 ExpressionType::COMPARE_LESSTHANOREQUALTO, ExpressionType::COMPARE_GREATERTHANOREQUALTO};
	const adacom_cmp cmps[6] = {ADACOM_CMP_EQ, ADACOM_CMP_NE, ADACOM_CMP_LT, ADACOM_CMP_GT, ADACOM_CMP_LE, ADACOM_CMP_GE};
	for (int k = 0; k < 6; k++) {
		auto st = ColumnData::BuildFilterStages(ConstantFilter(types[k], -7 + k));
		CHECK(st.size() == 1 && st[0].conj == ADACOM_CONJ_AND && st[0].preds.size() == 1);
		CHECK(st[0].preds[0].kind == ADACOM_PRED_COMPARE && st[0].preds[0].op == (int32_t)cmps[k] && st[0].preds[0].constant == -7 + k);
	}
	{
		auto st = ColumnData::BuildFilterStages(IsNullFilter());
		CHECK(st.size() == 1 && st[0].preds.size() == 1 && st[0].preds[0].kind == ADACOM_PRED_IS_NULL);
		st = ColumnData::BuildFilterStages(IsNotNullFilter());
		CHECK(st.size() == 1 && st[0].preds.size() == 1 && st[0].preds[0].kind == ADACOM_PRED_IS_NOT_NULL);
	}
	// BETWEEN with a nested AND: one call, three children
	{
		ConjunctionAndFilter f;
		f.child_filters.push_back(Cmp(ExpressionType::COMPARE_GREATERTHANOREQUALTO, 3));
		auto inner = new ConjunctionAndFilter();
		inner->child_filters.push_back(Cmp(ExpressionType::COMPARE_LESSTHAN, 90));
		inner->child_filters.push_back(unique_ptr<TableFilter>(new IsNotNullFilter()));
		f.child_filters.push_back(unique_ptr<TableFilter>(inner));
		auto st = ColumnData::BuildFilterStages(f);
		CHECK(st.size() == 1 && st[0].conj == ADACOM_CONJ_AND && st[0].preds.size() == 3);
	}
	// OR of leaves (nested ORs flatten): one OR call
	{
		ConjunctionOrFilter f;
		f.child_filters.push_back(Cmp(ExpressionType::COMPARE_EQUAL, 1));
		auto inner = new ConjunctionOrFilter();
		inner->child_filters.push_back(Cmp(ExpressionType::COMPARE_EQUAL, 2));
		inner->child_filters.push_back(unique_ptr<TableFilter>(new IsNullFilter()));
		f.child_filters.push_back(unique_ptr<TableFilter>(inner));
		auto st = ColumnData::BuildFilterStages(f);
		CHECK(st.size() == 1 && st[0].conj == ADACOM_CONJ_OR && st[0].preds.size() == 3);
	}
	// x >= 3 AND x IN (1, 2, 5): the leaves first, then the OR group refines their selection
	{
		ConjunctionAndFilter f;
		f.child_filters.push_back(Cmp(ExpressionType::COMPARE_GREATERTHANOREQUALTO, 3));
		auto in_list = new ConjunctionOrFilter();
		for (int64_t c : {1, 2, 5}) {
			in_list->child_filters.push_back(Cmp(ExpressionType::COMPARE_EQUAL, c));
		}
		f.child_filters.push_back(unique_ptr<TableFilter>(in_list));
		auto st = ColumnData::BuildFilterStages(f);
		CHECK(st.size() == 2);
		CHECK(st[0].conj == ADACOM_CONJ_AND && st[0].preds.size() == 1 && st[0].preds[0].op == (int32_t)ADACOM_CMP_GE);
		CHECK(st[1].conj == ADACOM_CONJ_OR && st[1].preds.size() == 3);
	}
	// only OR groups under the AND: no empty leading call
	{
		ConjunctionAndFilter f;
		for (int g = 0; g < 2; g++) {
			auto group = new ConjunctionOrFilter();
			group->child_filters.push_back(Cmp(ExpressionType::COMPARE_LESSTHAN, 10 * g));
			group->child_filters.push_back(Cmp(ExpressionType::COMPARE_GREATERTHAN, 100 * g));
			f.child_filters.push_back(unique_ptr<TableFilter>(group));
		}
		auto st = ColumnData::BuildFilterStages(f);
		CHECK(st.size() == 2 && st[0].conj == ADACOM_CONJ_OR && st[1].conj == ADACOM_CONJ_OR);
		CHECK(st[0].preds.size() == 2 && st[1].preds.size() == 2);
	}
	// an AND group under an OR is refused; so are more children than one call takes
	{
		ConjunctionOrFilter f;
		f.child_filters.push_back(unique_ptr<TableFilter>(new ConjunctionAndFilter()));
		bool threw = false;
		try {
			ColumnData::BuildFilterStages(f);
		} catch (const InternalException &) {
			threw = true;
		}
		CHECK(threw);
		ConjunctionAndFilter wide;
		for (int k = 0; k < ADACOM_MAX_PREDS + 1; k++) {
			wide.child_filters.push_back(Cmp(ExpressionType::COMPARE_NOTEQUAL, k));
		}
		threw = false;
		try {
			ColumnData::BuildFilterStages(wide);
		} catch (const InternalException &) {
			threw = true;
		}
		CHECK(threw);
	}
	printf("FILTER_STAGE_TESTS_OK\n");
	return 0;
}
