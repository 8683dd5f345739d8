"""Synthetic query-stream generator (stand-in for TPC `dsqgen -dialect spark`, absent here).

Writes a `query_<stream>.sql` in the grammar the reference's `spark.tpl` dialect defines
(/root/reference/nds/tpcds-gen/patches/templates.patch:468-472):

    -- start query <n> in stream <s> using template query<k>.tpl
    <sql> ;
    -- end query <n> in stream <s> using template query<k>.tpl

so that `gen_sql_from_stream` (/root/reference/nds/nds_power.py:50-77) splits it unchanged.
The statement texts below are this repo's own rendering of the star-join query *shapes* listed
in SURVEY.md section 8(a-2) (q3, q7, q19, q42, q52, q55); the TPC templates themselves are not
in the reference.  Substitution parameters are drawn from a seeded RNG, like dsqgen's RNGSEED.
"""
import numpy as np

from . import datagen

SUBSET = ['query3', 'query7', 'query19', 'query42', 'query52', 'query55']


def _params(seed, stream):
    r = np.random.default_rng([seed, 4242, stream])
    return {
        'MANUFACT': int(r.integers(1, 1001)),
        'MONTH': int(r.integers(11, 13)),
        'YEAR': int(r.integers(1998, 2003)),
        'MANAGER': int(r.integers(1, 101)),
        'GEN': ['M', 'F'][int(r.integers(0, 2))],
        'MS': datagen.MARITAL[int(r.integers(0, 5))],
        'ES': datagen.EDUCATION[int(r.integers(0, 7))],
        # query98 (drawn after the six-query stream's parameters, so those keep their values)
        'CAT1': 'Sports', 'CAT2': 'Books', 'CAT3': 'Home', 'SDATE': '1998-01-05',
    }


TEMPLATES = {
    'query3': """select dt.d_year, item.i_brand_id brand_id, item.i_brand brand, sum(ss_ext_sales_price) sum_agg
 from date_dim dt, store_sales, item
 where dt.d_date_sk = store_sales.ss_sold_date_sk
   and store_sales.ss_item_sk = item.i_item_sk
   and item.i_manufact_id = {MANUFACT}
   and dt.d_moy = {MONTH}
 group by dt.d_year, item.i_brand, item.i_brand_id
 order by dt.d_year, sum_agg desc, brand_id
 LIMIT 100""",
    'query7': """select i_item_id,
        avg(ss_quantity) agg1,
        avg(ss_list_price) agg2,
        avg(ss_coupon_amt) agg3,
        avg(ss_sales_price) agg4
 from store_sales, customer_demographics, date_dim, item, promotion
 where ss_sold_date_sk = d_date_sk and
       ss_item_sk = i_item_sk and
       ss_cdemo_sk = cd_demo_sk and
       ss_promo_sk = p_promo_sk and
       cd_gender = '{GEN}' and
       cd_marital_status = '{MS}' and
       cd_education_status = '{ES}' and
       (p_channel_email = 'N' or p_channel_event = 'N') and
       d_year = {YEAR}
 group by i_item_id
 order by i_item_id
 LIMIT 100""",
    'query19': """select i_brand_id brand_id, i_brand brand, i_manufact_id, i_manufact,
        sum(ss_ext_sales_price) ext_price
 from date_dim, store_sales, item, customer, customer_address, store
 where d_date_sk = ss_sold_date_sk
   and ss_item_sk = i_item_sk
   and i_manager_id = {MANAGER}
   and d_moy = {MONTH}
   and d_year = {YEAR}
   and ss_customer_sk = c_customer_sk
   and c_current_addr_sk = ca_address_sk
   and substr(ca_zip,1,5) <> substr(s_zip,1,5)
   and ss_store_sk = s_store_sk
 group by i_brand, i_brand_id, i_manufact_id, i_manufact
 order by ext_price desc, i_brand, i_brand_id, i_manufact_id, i_manufact
 LIMIT 100""",
    'query42': """select dt.d_year, item.i_category_id, item.i_category, sum(ss_ext_sales_price)
 from date_dim dt, store_sales, item
 where dt.d_date_sk = store_sales.ss_sold_date_sk
   and store_sales.ss_item_sk = item.i_item_sk
   and item.i_manager_id = 1
   and dt.d_moy = {MONTH}
   and dt.d_year = {YEAR}
 group by dt.d_year, item.i_category_id, item.i_category
 order by sum(ss_ext_sales_price) desc, dt.d_year, item.i_category_id, item.i_category
 LIMIT 100""",
    'query52': """select dt.d_year, item.i_brand_id brand_id, item.i_brand brand, sum(ss_ext_sales_price) ext_price
 from date_dim dt, store_sales, item
 where dt.d_date_sk = store_sales.ss_sold_date_sk
   and store_sales.ss_item_sk = item.i_item_sk
   and item.i_manager_id = 1
   and dt.d_moy = {MONTH}
   and dt.d_year = {YEAR}
 group by dt.d_year, item.i_brand, item.i_brand_id
 order by dt.d_year, ext_price desc, brand_id
 LIMIT 100""",
    'query55': """select i_brand_id brand_id, i_brand brand, sum(ss_ext_sales_price) ext_price
 from date_dim, store_sales, item
 where d_date_sk = ss_sold_date_sk
   and ss_item_sk = i_item_sk
   and i_manager_id = {MANAGER}
   and d_moy = {MONTH}
   and d_year = {YEAR}
 group by i_brand, i_brand_id
 order by ext_price desc, i_brand_id
 LIMIT 100""",
}


# Statements beyond the six-query bench stream that the planner and the engine also run (tests; not part of the metric)
EXTRA = ['query98']
TEMPLATES['query98'] = """select i_item_id, i_item_desc, i_category, i_class, i_current_price,
       sum(ss_ext_sales_price) as itemrevenue,
       sum(ss_ext_sales_price)*100/sum(sum(ss_ext_sales_price)) over (partition by i_class) as revenueratio
 from store_sales, item, date_dim
 where ss_item_sk = i_item_sk
   and i_category in ('{CAT1}', '{CAT2}', '{CAT3}')
   and ss_sold_date_sk = d_date_sk
   and d_date between cast('{SDATE}' as date) and (cast('{SDATE}' as date) + interval 30 days)
 group by i_item_id, i_item_desc, i_category, i_class, i_current_price
 order by i_category, i_class, i_item_id, i_item_desc, revenueratio"""


def render(template_name, seed=datagen.SEED, stream=0, overrides=None):
    p = _params(seed, stream)
    if overrides:
        p.update(overrides)
    return TEMPLATES[template_name].format(**p)


def generate_stream(path, seed=datagen.SEED, stream=0, queries=None, overrides=None):
    """Write one query stream file; returns the list of query names in stream order."""
    queries = list(queries or SUBSET)
    with open(path, 'w') as f:
        for i, q in enumerate(queries):
            head = 'query %d in stream %d using template %s.tpl' % (i + 1, stream, q)
            f.write('-- start %s\n%s\n;\n-- end %s\n' % (head, render(q, seed, stream, overrides), head))
    return queries
