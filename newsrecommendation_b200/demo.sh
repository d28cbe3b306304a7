#!/bin/bash
# demo.sh train | test <ckpt> | embeddings — the reference's src/demo.sh flow (same variables and flags)
# on the sm_100a path: python -m newsrecommendation_b200.main instead of src/main.py.

mode=$1
nGPU=${nGPU:-1}
enable_gpu=True
model_dir=${model_dir:-'../model/NAML'}
model=${model:-'NAML'}
use_category=True
use_subcategory=True
train_data_dir=${train_data_dir:-'../data/MINDsmall_train'}
test_data_dir=${test_data_dir:-'../data/MINDsmall_dev'}
freeze_embedding=True
prepare=${prepare:-False}
here=$(cd "$(dirname "$0")/.." && pwd)

if [ "${mode}" == train ]
then
 epochs=${epochs:-5}
 batch_size=${batch_size:-32}
 lr=0.0003
 user_log_mask=False
 PYTHONPATH=${here} python -u -m newsrecommendation_b200.main --mode train --model_dir ${model_dir} --batch_size ${batch_size} \
 --epochs ${epochs} --model ${model} --lr ${lr} --user_log_mask ${user_log_mask} --prepare ${prepare} --nGPU ${nGPU} \
 --enable_gpu ${enable_gpu} --use_category ${use_category} --use_subcategory ${use_subcategory} \
 --train_data_dir ${train_data_dir} --freeze_embedding ${freeze_embedding} "${@:2}"
elif [ "${mode}" == test ]
then
 user_log_mask=True
 batch_size=${batch_size:-128}
 load_ckpt_name=$2
 PYTHONPATH=${here} python -u -m newsrecommendation_b200.main --mode test --model_dir ${model_dir} --batch_size ${batch_size} \
 --user_log_mask ${user_log_mask} --load_ckpt_name ${load_ckpt_name} --model ${model} --prepare ${prepare} --nGPU ${nGPU} \
 --enable_gpu ${enable_gpu} --use_category ${use_category} --use_subcategory ${use_subcategory} \
 --test_data_dir ${test_data_dir} --freeze_embedding ${freeze_embedding} "${@:3}"
elif [ "${mode}" == embeddings ]
then
 PYTHONPATH=${here} python -u -m newsrecommendation_b200.main --mode read_embeddings \
 --train_data_dir ${train_data_dir} --test_data_dir ${test_data_dir}
else
 echo "Please select train or test mode."
fi
